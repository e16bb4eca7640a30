// mmvq.cuh -- decode GEMV / skinny GEMM over GGUF block formats for B200 (sm_100a), v2.
//
// Replaces the reference's two-kernel path  quantize_q8_1 -> mul_mat_vec_q
// (ggml-cuda/mmvq.cu:491-561, :140-224; quantize.cu:4-49) with ONE warp-specialised kernel:
//
//   warp 0      producer + epilogue.  Streams whole weight rows HBM -> shared memory with 1-D TMA
//               bulk copies (cp.async.bulk, mbarrier complete_tx, L2 evict-first) through a ring
//               of stages of R rows; it starts BEFORE the programmatic-dependent-launch wait, so
//               the first stages of this GEMV are in flight while the previous kernel drains.
//               When the consumers release a stage it adds their partial sums in a fixed order
//               and stores dst.
//   warps 1..   consumers, NG groups of CW warps; group j owns stages j, j+NG, ...  They first
//               quantise the M activation columns ONCE per CTA into shared memory (fused q8_1
//               prologue, mmvq_units.cuh), then every thread takes 256-weight items
//               (row, b) of its stage: lanes are consecutive items, so a row is a run of
//               consecutive lanes and the row sum is a segmented xor-shuffle over G lanes
//               (G = largest power of two dividing the items per row, <= 32) followed by
//               BPR/G partials per row in shared memory.
//
// Weights are read exactly once, in 16-byte-aligned bulk transactions.  The kernel is sized so
// that two CTAs (this GEMV's, or this one and the next kernel's) fit one SM: <= 113 KB of shared
// memory and 288 threads, which is what lets consecutive kernels of a decode step overlap under
// PDL instead of leaving HBM idle at every kernel boundary.
#pragma once
#include "mmvq_units.cuh"

namespace b200 {

struct mmvq_params {
    const uint8_t * W;       // weight rows
    int64_t nb1;             // bytes between rows
    const float * X;         // activations, column m at X + m*xs
    const uint8_t * Xq;      // optional: activations already quantised (mmvq_units.cuh layout, column stride BPR*400)
    int64_t xs;
    float * dst;             // dst[m*ds + row]
    int64_t ds;
    int N, K;
    // batch/expert channels: blockIdx.y = y, blockIdx.z = z (see b200_mmvq_args)
    int64_t w_sy, w_sz;      // bytes
    int w_div_y, w_div_z;
    int64_t x_sy, x_sz;      // floats
    int x_mod_y;
    int64_t d_sy, d_sz;      // floats
    const int32_t * ids; int64_t ids_sz;
    int64_t xq_sy, xq_sz;    // bytes
    // launch geometry
    int R;                   // rows per stage
    int nstages;
    int pitch;               // bytes between rows inside a stage (== row bytes in contiguous mode)
    int stage_bytes;         // multiple of 16, >= R*pitch + 16
    int contig;              // 1: one bulk copy per stage (rows back to back, 16 B aligned)
    int CW, NG;              // consumer warps per group, groups
    int BPR;                 // items (256 weights) per row = ceil(K/256)
    int G;                   // lanes per reduction segment (power of two dividing BPR)
    int rows_q, rows_rem;    // N / gridDim.x, N % gridDim.x
    int act_bytes;           // shared bytes of the activation area (M columns)
    int vec_ok;              // X columns 16-byte aligned -> float4 loads
    int32_t * dbg_sumi;      // parity hook: when non-null, int32 block sums [M][N][K/32] are stored here
    // group mode (b200_mul_mat_vec_q_group): up to 3 weight matrices of one type and K read the same activations in one
    // launch; CTAs [cta0[g], cta0[g+1]) work on matrix g with their own balanced row split
    int ngroup;
    const uint8_t * Wg[3]; float * Dg[3]; int64_t dsg[3];
    int cta0[4], rows_qg[3], rows_remg[3];
    // trailing L2 prefetch of the next matmul's weights (b200_mmvq_args.pf_*)
    int pf_n; const uint8_t * pf_ptr[3]; int64_t pf_bytes[3];
    hs_params hs;            // row-split hand-over folded into the launch (common.cuh); needs Xq
};

template <int TYPE, int M, bool A16, bool DBG>
__global__ void __launch_bounds__(288, (M <= 4 ? 2 : 1)) mmvq_kernel(const mmvq_params p) {
    constexpr int IB = item_traits<TYPE>::BYTES;
    extern __shared__ __align__(128) uint8_t smem_raw[];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int ncw = p.CW * p.NG;                   // consumer warps
    const int nb32 = p.K >> 5;
    const int nseg = p.BPR / p.G;
    const uint32_t acs = (uint32_t)p.BPR * ACT_ITEM;   // bytes between activation columns

    // shared memory: [stages][activations][partials][barriers]
    uint8_t * stages = smem_raw;
    uint8_t * act = smem_raw + (size_t)p.nstages * p.stage_bytes;
    float * partials = reinterpret_cast<float *>(act + p.act_bytes);          // [nstages][R][nseg][M]
    const int part_stage = p.R * nseg * M;
    uint64_t * bars = reinterpret_cast<uint64_t *>(partials + (((size_t)p.nstages * part_stage + 1) & ~(size_t)1));
    const uint32_t full0 = smem_u32(bars), empty0 = full0 + 8 * p.nstages, actbar = empty0 + 8 * p.nstages;

    // channel (batch / expert) addressing
    const uint8_t * Wc = p.W;
    const float * Xc = p.X;
    const uint8_t * Xqc = p.Xq;
    float * Dc = p.dst;
    if (gridDim.y > 1 || gridDim.z > 1 || p.ids) {
        const int y = blockIdx.y, z = blockIdx.z;
        if (p.ids) { pdl_wait(); Wc += (int64_t)p.ids[z * p.ids_sz + y] * p.w_sy; }      // the expert ids come from the previous kernels
        else Wc += (int64_t)(y / p.w_div_y) * p.w_sy + (int64_t)(z / p.w_div_z) * p.w_sz;
        const int xy = y % p.x_mod_y;
        if (Xc) Xc += (int64_t)xy * p.x_sy + (int64_t)z * p.x_sz;
        if (Xqc) Xqc += (int64_t)xy * p.xq_sy + (int64_t)z * p.xq_sz;
        Dc += (int64_t)y * p.d_sy + (int64_t)z * p.d_sz;
    }

    // balanced contiguous row range of this CTA: the first rows_rem CTAs take one extra row
    int cta = blockIdx.x, rows_q = p.rows_q, rows_rem = p.rows_rem;
    int64_t ds = p.ds;
    if (p.ngroup) {
        int g = 0;
        while (g + 1 < p.ngroup && cta >= p.cta0[g + 1]) ++g;
        cta -= p.cta0[g]; Wc = p.Wg[g]; Dc = p.Dg[g]; ds = p.dsg[g]; rows_q = p.rows_qg[g]; rows_rem = p.rows_remg[g];
    }
    const int row_lo = cta * rows_q + min(cta, rows_rem);
    const int nrows = rows_q + (cta < rows_rem ? 1 : 0);
    const int row_hi = row_lo + nrows;
    const int ngroups = (nrows + p.R - 1) / p.R;
    const int rb = (p.K / fmt<TYPE>::BE) * fmt<TYPE>::BB;   // payload bytes per row

    if (tid == 0) {
        for (int s = 0; s < p.nstages; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, p.CW); }
        mbar_init(actbar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    pdl_trigger();            // dependents may start their own prologue (weight prefetch) now

    if (warp == 0) {
        // ================= producer + epilogue warp =================
        const uint64_t pol = policy_evict_first();
        auto issue = [&](int g) {   // lane 0 only
            const int slot = g % p.nstages;
            const int r0 = row_lo + g * p.R, nr = min(p.R, row_hi - r0);
            const uint32_t dst = smem_u32(stages + (size_t)slot * p.stage_bytes), bar = full0 + 8 * slot;
            if (p.contig) {
                const uint32_t bytes = (uint32_t)nr * rb;
                mbar_expect_tx(bar, bytes);
                bulk_g2s_stream(dst, Wc + (int64_t)r0 * p.nb1, bytes, bar, pol);
            } else {
                uint32_t total = 0;
                for (int r = 0; r < nr; ++r) {
                    const uint64_t ga = (uint64_t)(Wc + (int64_t)(r0 + r) * p.nb1);
                    total += (uint32_t)(((ga & 15) + rb + 15) & ~15ull);
                }
                mbar_expect_tx(bar, total);
                for (int r = 0; r < nr; ++r) {
                    const uint64_t ga = (uint64_t)(Wc + (int64_t)(r0 + r) * p.nb1);
                    const uint32_t bytes = (uint32_t)(((ga & 15) + rb + 15) & ~15ull);
                    bulk_g2s_stream(dst + r * p.pitch, (const void *)(ga & ~15ull), bytes, bar, pol);
                }
            }
        };
        // once this CTA's last weight load is queued: its share of the next matmul's weights -> L2 (hint only)
        auto prefetch_next = [&]() {   // lane 0 only
            const long long nct = (long long)gridDim.x * gridDim.y * gridDim.z;
            const long long cta = blockIdx.x + (long long)gridDim.x * (blockIdx.y + (long long)gridDim.y * blockIdx.z);
            for (int r = 0; r < p.pf_n; ++r) {
                const long long per = ((p.pf_bytes[r] + nct - 1) / nct + 15) & ~15ll;
                long long lo = cta * per;
                const long long hi = min(p.pf_bytes[r] & ~15ll, lo + per);
                for (; lo < hi; lo += 16384) {
                    const unsigned size = (unsigned)min(16384ll, hi - lo);
                    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(p.pf_ptr[r] + lo), "r"(size) : "memory");
                }
            }
        };
        // weights do not depend on the previous kernel: start streaming before the PDL wait
        if (lane == 0) {
            for (int g = 0; g < min(ngroups, p.nstages); ++g) issue(g);
            if (p.pf_n && ngroups <= p.nstages) prefetch_next();
        }
        pdl_wait();   // dst may still be read by the previous grid; Xq is written by it
        if (lane == 0) hs_wait_acts(p.hs);      // row split, peer side: the main device's broadcast has landed
        if (p.Xq && lane == 0) {
            const uint8_t * xq = Xqc;
            mbar_expect_tx(actbar, (uint32_t)M * acs);
            for (int m = 0; m < M; ++m) bulk_g2s(smem_u32(act) + m * acs, xq + (size_t)m * acs, acs, actbar);
        }
        for (int g = 0; g < ngroups; ++g) {
            const int slot = g % p.nstages;
            mbar_wait(empty0 + 8 * slot, (g / p.nstages) & 1);
            const int r0 = row_lo + g * p.R, nr = min(p.R, row_hi - r0);
            const float * ps = partials + (size_t)slot * part_stage;
            for (int o = lane; o < nr * M; o += 32) {
                const int r = o / M, m = o - r * M;
                float s = 0.0f;
                for (int sg = 0; sg < nseg; ++sg) s += ps[(r * nseg + sg) * M + m];
                Dc[(int64_t)m * ds + r0 + r] = s;
            }
            __syncwarp();
            if (lane == 0 && g + p.nstages < ngroups) {
                issue(g + p.nstages);
                if (p.pf_n && g + p.nstages == ngroups - 1) prefetch_next();
            }
        }
        if (p.hs.counter) {                    // row split: the last CTA of the launch lowers / raises / awaits the hand-over flags
            __syncwarp();
            if (lane == 0) hs_finish(p.hs, gridDim.x * gridDim.y * gridDim.z);
        }
    } else {
        // ================= consumer warps =================
        // all eight take part in the activation prologue; CW*NG of them go on to the dot products
        const int cw = warp - 1, grp = cw / p.CW, wig = cw - grp * p.CW;   // warp inside its group
        const int T = 32 * p.CW;
        if (p.Xq) {
            if (cw >= ncw) return;
            mbar_wait(actbar, 0);
        } else {
            pdl_wait();   // activations come from the previous kernel
            quantize_acts(Xc, p.xs, M, nb32, p.vec_ok, act, acs, tid - 32, 256);
            asm volatile("bar.sync 1, 256;" ::: "memory");
            if (cw >= ncw) return;
        }
        const uint32_t act0 = smem_u32(act);
        const int dT_row = T / p.BPR, dT_b = T - dT_row * p.BPR;       // item-index step per pass
        for (int g = grp; g < ngroups; g += p.NG) {
            const int slot = g % p.nstages;
            const int r0 = row_lo + g * p.R, nr = min(p.R, row_hi - r0);
            const int nitems = nr * p.BPR;
            mbar_wait(full0 + 8 * slot, (g / p.nstages) & 1);
            const uint32_t sbase = smem_u32(stages + (size_t)slot * p.stage_bytes);
            float * pw = partials + (size_t)slot * part_stage;
            int i = wig * 32 + lane;
            int row = i / p.BPR, b = i - row * p.BPR;
            for (int i0 = wig * 32; i0 < nitems; i0 += T) {
                float acc[M];
#pragma unroll
                for (int m = 0; m < M; ++m) acc[m] = 0.0f;
                const bool valid = i < nitems;
                if (valid) {
                    uint32_t wp = sbase + (uint32_t)row * p.pitch + (uint32_t)b * IB;
                    if (!p.contig) wp += (uint32_t)((uint64_t)(Wc + (int64_t)(r0 + row) * p.nb1) & 15);
                    const int nblk = min(8, nb32 - 8 * b);
                    int32_t * dbg = DBG ? p.dbg_sumi + (int64_t)(r0 + row) * nb32 + 8 * b : nullptr;
                    item_dot<TYPE, M, A16, DBG>(wp, act0 + (uint32_t)b * ACT_ITEM, acs, nblk, acc, dbg, (int64_t)p.N * nb32);
                }
                // segmented row sum over G consecutive lanes (G is warp-uniform)
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    if (o < p.G) {
#pragma unroll
                        for (int m = 0; m < M; ++m) acc[m] += __shfl_xor_sync(0xffffffffu, acc[m], o);
                    }
                }
                if (valid && (lane & (p.G - 1)) == 0) {
                    float * dstp = pw + (row * nseg + b / p.G) * M;
#pragma unroll
                    for (int m = 0; m < M; ++m) dstp[m] = acc[m];
                }
                i += T; row += dT_row; b += dT_b;
                if (b >= p.BPR) { b -= p.BPR; ++row; }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty0 + 8 * slot);
        }
    }
}

}  // namespace b200
