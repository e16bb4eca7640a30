// mmvq.cuh -- decode GEMV / skinny GEMM over GGUF block formats for B200 (sm_100a).
//
// Replaces the reference's two-kernel path  quantize_q8_1 -> mul_mat_vec_q
// (ggml-cuda/mmvq.cu:491-561, :140-224; quantize.cu:4-49) with ONE warp-specialised kernel:
//
//   warp 0      producer + epilogue: streams whole weight rows HBM -> shared memory with 1-D TMA
//               bulk copies (cp.async.bulk, mbarrier complete_tx) through a ring of stages, and,
//               when all consumers have released a stage, adds their per-warp partials in a fixed
//               order and stores dst.
//   warps 1..   consumers, organised in NT teams of WPT warps.  A team covers one row: thread t of
//               the team owns units t, t+T, ... (T = 32*WPT) and keeps the q8_1-quantised
//               activations of those units in REGISTERS for the whole kernel (fused prologue).
//               Row r of a stage belongs to team r % NT.
//
// Weights are read exactly once, in 16-byte-aligned bulk transactions; activations are re-read
// (fp32, from L2) once per team per CTA.  No __syncthreads after setup: all hand-offs are mbarriers.
#pragma once
#include "mmvq_units.cuh"

namespace b200 {

struct mmvq_params {
    const uint8_t * W;       // weight rows
    int64_t nb1;             // bytes between rows
    const float * X;         // activations, column m at X + m*xs
    int64_t xs;
    float * dst;             // dst[m*ds + row]
    int64_t ds;
    int N, K;
    // batch/expert channels (blockIdx.y): per-channel offsets; ids selects the weight matrix
    int64_t w_ch_stride;     // bytes between weight matrices (experts / broadcast dims)
    int64_t x_ch_stride;     // floats
    int64_t d_ch_stride;     // floats
    const int32_t * ids;     // optional [nchannels]: expert index per channel (MUL_MAT_ID decode)
    int x_ch_mod;            // channel c reads X channel (c % x_ch_mod) (ne11 broadcast of MUL_MAT_ID)
    int w_ch_div;            // channel c reads W matrix (c / w_ch_div) when ids == nullptr (ne02 broadcast)
    // launch geometry
    int R;                   // rows per stage
    int nstages;
    int pitch;               // bytes between rows inside a stage (== row bytes in contiguous mode)
    int stage_bytes;         // R*pitch + 16 slack, multiple of 16
    int contig;              // 1: one bulk copy per stage (rows back to back, 16 B aligned)
    int WPT, NT;             // warps per team, teams
    int vec_ok;              // X columns 16-byte aligned -> float4 loads
    int32_t * dbg_sumi;      // parity hook: when non-null, int32 block sums [M][N][K/32] are accumulated here
};

// Multi-value warp reduction: NV values per lane (NV a power of two <= 32).  After the call lane l
// holds the full sum of value index reduce_index<NV>(l) in v[0].  NV-1 + (5-log2 NV) shuffles.
template <int NV> __device__ __forceinline__ void warp_reduce_multi(float (&v)[NV], int lane) {
    int off = 16;
#pragma unroll
    for (int n = NV; n > 1; n >>= 1, off >>= 1) {
        const bool up = lane & off;
#pragma unroll
        for (int i = 0; i < n / 2; ++i) {
            const float send = up ? v[i] : v[i + n / 2];
            const float keep = up ? v[i + n / 2] : v[i];
            v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
        }
    }
    for (; off > 0; off >>= 1) v[0] += __shfl_xor_sync(0xffffffffu, v[0], off);
}
template <int NV> __device__ __forceinline__ int reduce_index(int lane) {
    int idx = 0, off = 16;
#pragma unroll
    for (int n = NV; n > 1; n >>= 1, off >>= 1) if (lane & off) idx += n / 2;
    return idx;
}

constexpr int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }

// rows a team accumulates before one multi-value reduction
template <int M> struct row_batch { static constexpr int RB = M == 1 ? 4 : (M == 2 ? 2 : 1); };

template <int TYPE, int M, int CPT, bool A16>
__global__ void __launch_bounds__(544, 1) mmvq_kernel(const mmvq_params p) {
    constexpr int G = unit_traits<TYPE>::G;
    constexpr int RB = row_batch<M>::RB;
    constexpr int NV = next_pow2(RB * M);
    extern __shared__ __align__(128) uint8_t smem_raw[];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int U = p.K >> 5;                        // units per row
    const int T = 32 * p.WPT;                      // threads per team
    const int ncw = p.WPT * p.NT;                  // consumer warps
    const int rows_pt = p.R / p.NT;                // rows per team per stage (R is a multiple of NT*RB)

    // shared memory carve-up: [stages][partials][barriers]
    uint8_t * stages = smem_raw;
    float * partials = reinterpret_cast<float *>(smem_raw + (size_t)p.nstages * p.stage_bytes);   // [nstages][ncw][rows_pt*M]
    // mbarriers need 8-byte alignment: round the partials area up to an even number of floats
    uint64_t * bars = reinterpret_cast<uint64_t *>(partials + (((size_t)p.nstages * ncw * rows_pt * M + 1) & ~(size_t)1));
    const uint32_t full0 = smem_u32(bars), empty0 = full0 + 8 * p.nstages;
    const int part_stage = ncw * rows_pt * M;

    // channel (batch / expert) addressing
    const int ch = blockIdx.y;
    const int wsel = p.ids ? p.ids[ch] : ch / p.w_ch_div;
    const uint8_t * Wc = p.W + (int64_t)wsel * p.w_ch_stride;
    const float * Xc = p.X + (int64_t)(ch % p.x_ch_mod) * p.x_ch_stride;
    float * Dc = p.dst + (int64_t)ch * p.d_ch_stride;

    // balanced contiguous row range of this CTA
    const int nct = gridDim.x, cta = blockIdx.x;
    const int row_lo = (int)(((int64_t)p.N * cta) / nct), row_hi = (int)(((int64_t)p.N * (cta + 1)) / nct);
    const int nrows = row_hi - row_lo;
    const int ngroups = (nrows + p.R - 1) / p.R;
    const int rb = (p.K / fmt<TYPE>::BE) * fmt<TYPE>::BB;   // payload bytes per row

    if (tid == 0) {
        for (int s = 0; s < p.nstages; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, ncw); }
        mbar_fence_init();
    }
    __syncthreads();

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
        // weights do not depend on the previous kernel: start streaming before the PDL wait
        if (lane == 0) for (int g = 0; g < min(ngroups, p.nstages); ++g) issue(g);
        pdl_wait();   // dst may still be read by the previous grid
        for (int g = 0; g < ngroups; ++g) {
            const int slot = g % p.nstages;
            mbar_wait(empty0 + 8 * slot, (g / p.nstages) & 1);
            const int r0 = row_lo + g * p.R, nr = min(p.R, row_hi - r0);
            const float * ps = partials + (size_t)slot * part_stage;
            for (int o = lane; o < nr * M; o += 32) {
                const int r = o / M, m = o - r * M;
                const int team = r % p.NT, rt = r / p.NT;       // row rt of that team's share
                float s = 0.0f;
                for (int w = 0; w < p.WPT; ++w) s += ps[((team * p.WPT + w) * rows_pt + rt) * M + m];
                Dc[(int64_t)m * p.ds + r0 + r] = s;
            }
            __syncwarp();
            if (lane == 0 && g + p.nstages < ngroups) issue(g + p.nstages);
        }
        pdl_trigger();
    } else if (warp <= ncw) {
        // ================= consumer warps =================
        const int cw = warp - 1, team = cw / p.WPT, tt = (cw % p.WPT) * 32 + lane;   // thread index inside team
        pdl_wait();   // activations come from the previous kernel
        act_regs<G> a[CPT][M];
#pragma unroll
        for (int c = 0; c < CPT; ++c) {
            const int u = tt + c * T;
#pragma unroll
            for (int m = 0; m < M; ++m) quantize_unit<TYPE>(Xc + (int64_t)m * p.xs, u, u < U, p.vec_ok, a[c][m]);
        }
        const int ridx = reduce_index<NV>(lane);
        for (int g = 0; g < ngroups; ++g) {
            const int slot = g % p.nstages;
            const int r0 = row_lo + g * p.R, nr = min(p.R, row_hi - r0);
            mbar_wait(full0 + 8 * slot, (g / p.nstages) & 1);
            const uint32_t sbase = smem_u32(stages + (size_t)slot * p.stage_bytes);
            float * pw = partials + (size_t)slot * part_stage + (size_t)cw * rows_pt * M;
            for (int rt0 = 0; rt0 < rows_pt; rt0 += RB) {
                float v[NV];
#pragma unroll
                for (int i = 0; i < NV; ++i) v[i] = 0.0f;
#pragma unroll
                for (int rbi = 0; rbi < RB; ++rbi) {
                    const int r = (rt0 + rbi) * p.NT + team;     // row inside the stage
                    if (r < nr) {
                        uint32_t rowp = sbase + (uint32_t)r * p.pitch;
                        if (!p.contig) rowp += (uint32_t)((uint64_t)(Wc + (int64_t)(r0 + r) * p.nb1) & 15);
                        float acc[M];
#pragma unroll
                        for (int m = 0; m < M; ++m) acc[m] = 0.0f;
#pragma unroll
                        for (int c = 0; c < CPT; ++c) {
                            const int u = tt + c * T;
                            if (u < U) {
                                if (p.dbg_sumi) {
                                    unit_ints<G> ints[M];
                                    unit_dot<TYPE, M, A16>(rowp, u, a[c], acc, ints);
                                    for (int m = 0; m < M; ++m) for (int gg = 0; gg < G; ++gg)
                                        atomicAdd(p.dbg_sumi + ((int64_t)m * p.N + r0 + r) * U + unit_group_elem0<TYPE>(u, gg) / 32, ints[m].dot[gg]);
                                } else {
                                    unit_dot<TYPE, M, A16>(rowp, u, a[c], acc);
                                }
                            }
                        }
#pragma unroll
                        for (int m = 0; m < M; ++m) v[rbi * M + m] = acc[m];
                    }
                }
                warp_reduce_multi<NV>(v, lane);
                // lane l holds value index ridx = rbi*M + m (lanes with equal ridx hold the same sum)
                if (ridx < RB * M && (lane & (32 / NV - 1)) == 0) {
                    const int rbi = ridx / M, m = ridx - rbi * M;
                    if (rt0 + rbi < rows_pt) pw[(rt0 + rbi) * M + m] = v[0];
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty0 + 8 * slot);
        }
    }
}

}  // namespace b200
