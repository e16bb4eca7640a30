// dchain.cu -- the mat-vec CHAIN of a decode step as ONE persistent kernel (b200_decode_chain).
//
// A batch-1 decode step of a Llama layer is, between two attentions,
//     wo mat-vec (+ residual) -> RMS_NORM * w -> gate / up mat-vecs -> SILU(gate) * up -> down mat-vec (+ residual)
//     -> RMS_NORM * w (next layer) -> q / k / v mat-vecs (+ rope, KV-cache store)        [or: -> lm_head after the last layer]
// The reference runs one kernel (or two) per node (ggml-cuda/ggml-cuda.cu:2150-2400 ggml_cuda_compute_forward, mmvq.cu:491-561,
// norm.cu, unary.cu, binbcast.cu, rope.cu, cpy.cu), round 1 of this backend 4-5 launches with the element-wise nodes riding in
// the mat-vecs' prologues / epilogues (mmvq_r.cuh).  Every dependent launch costs ~3.3 us of serial latency whatever it computes
// (launch, first byte, last CTA, grid completion; profiles/r01_ktrace_tg_v6.txt), the norm prologue 6 us more: 0.35 of the HBM
// roofline for the whole step with a kernel body that reaches 0.99 on its own.
//
// Here the CTAs stay resident across the whole chain (2 per SM, same shape as mmvqr_kernel: warp 0 = TMA producer + epilogue, 8
// consumer warps with register-resident activations):
//   * warp 0 streams weight rows through the shared-memory ring in chain order and NEVER waits for a phase boundary: while the
//     consumers sit at a grid barrier or normalise the residual stream, the rows of the next matrices keep arriving (the ring of
//     all CTAs holds ~21 MB, 3 us of HBM time), so HBM idles only when a stall outlasts the ring;
//   * phases hand over through counters in L2 (one atomic per CTA + acquire polling, ~1 us) instead of kernel boundaries;
//   * the arithmetic of every step is the code of the stand-alone kernels (half_dot / rope_pair / sumsq4 / quant_block_store, the
//     same partial-sum orders), so every tensor of the graph gets the bits the node-by-node path gives it
//     (tests/test_gpu_chain.py: chain == sequence of stand-alone launches, bit for bit).
// All CTAs must be co-resident (they wait for each other): the host sizes the grid from the occupancy query and the kernel is
// launched cooperatively sized; every spin is bounded (error word + carry on, no trap).
#include "mmvq_r.cuh"
#include "../../include/b200_ops.h"

#include <algorithm>
#include <cstdio>
#include <cstdlib>

namespace b200 {

constexpr int CH_MAXPH = 5;       // wo, gate/up, down, tail 1, tail 2
constexpr int CH_NBAR = 4;        // grid barriers per launch, always all four (the counters reset each other in a ring)
constexpr int CH_PART = 64;       // floats of row partials per stage (R * NW <= 64)

struct ch_gemv {
    const uint8_t * W[3]; float * D[3];
    int cta0[4]; int rows_q[3], rows_rem[3];             // CTAs [cta0[g], cta0[g+1]) take matrix g; rows in units of `unit`
    int ng, type, K, BPR, TPR, lgTPR, rb, R;
    int epi;                                             // 0 store, 1 store + residual sum, 2 rope / KV store
    const float * resid; float * dsum;
    int rmode[3]; float * rout[3]; char * rcell[3]; char * const * rslot[3]; int rctype[3], rcesz[3], rhd[3];
    const int32_t * rpos; const float * rff; rope_dev rp;
    int pre;                                             // activations: 0 copy the quantised column xq, 1 rms_norm(x) * w (norms[norm]), 2 keep the previous phase's
    const uint8_t * xq; int norm;
    int wait_bar, arrive_bar;                            // grid barrier passed before `pre` / fed by the completion of this phase (-1: none)
    int silu_before;                                     // the distributed SILU * MUL + quantisation step (and its barriers) comes first
};
struct ch_norm { const float * x; const float * w; float * y; float * out; float eps; int K; };
struct chain_params {
    ch_gemv ph[CH_MAXPH]; int nph;
    ch_norm norms[2];
    const float * sg; const float * su; float * sact; float * sout; uint8_t * sxq; int sn, sop, silu_wait, silu_arrive;
    uint32_t * bars;                                     // CH_NBAR counters, 128 bytes apart, zero between launches
    uint32_t * error; unsigned long long timeout_ns;
    int nstages, stage_bytes, act_bytes, wn_bytes, rsd_rows, csn_rows;
    unsigned long long * trace;                          // debug: globaltimer stamps of the first and the last CTA, 64 slots each (tools/chain_micro.py)
};
__device__ __forceinline__ void ch_stamp(const chain_params & p, int ev) {      // call from ONE thread of the warp
    if (!p.trace || (blockIdx.x != 0 && blockIdx.x != gridDim.x - 1)) return;
    unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.trace[(blockIdx.x ? 64 : 0) + ev] = t;
}

__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t * p) {
    uint32_t v; asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}
// one thread per CTA.  Counter k completes at gridDim.x arrivals; whoever completes it zeroes counter k-1 (every CTA has left
// that one: it arrived here after passing it), so the counters are back at zero for the next launch without a memset.
__device__ __forceinline__ void grid_arrive(const chain_params & p, int k) {
    __threadfence();
    const unsigned old = atomicAdd(p.bars + 32 * k, 1u);
    if (old == gridDim.x - 1) *(volatile uint32_t *)(p.bars + 32 * ((k + CH_NBAR - 1) % CH_NBAR)) = 0u;
}
__device__ __forceinline__ void grid_wait(const chain_params & p, int k) {
    const uint32_t * c = p.bars + 32 * k;
    unsigned long long t0 = 0;
    int n = 0;
    while (ld_acquire_gpu(c) < gridDim.x) {
        if ((++n & 1023) == 0 && p.timeout_ns) {
            unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t0 == 0) t0 = t;
            else if (t - t0 > p.timeout_ns) { if (p.error) *(volatile uint32_t *)p.error = 2u; break; }
        }
    }
}

struct ph_cta { int g, row_lo, nrows, nst; };
__device__ __forceinline__ ph_cta phase_rows(const ch_gemv & q) {
    ph_cta c; c.g = 0; c.row_lo = 0; c.nrows = 0; c.nst = 0;
    int cta = blockIdx.x;
    if (cta >= q.cta0[q.ng]) return c;
    while (c.g + 1 < q.ng && cta >= q.cta0[c.g + 1]) ++c.g;
    cta -= q.cta0[c.g];
    const int unit = q.epi == 2 ? 2 : 1;
    c.row_lo = unit * (cta * q.rows_q[c.g] + min(cta, q.rows_rem[c.g]));
    c.nrows = unit * (q.rows_q[c.g] + (cta < q.rows_rem[c.g] ? 1 : 0));
    c.nst = (c.nrows + q.R - 1) / q.R;
    return c;
}

__device__ __forceinline__ float chain_act(int op, float x) {      // the expressions of fused.cu act_fn
    if (op == B200_UNARY_SILU) return x / (1.0f + expf(-x));
    const float c = 0.044715f, s = 0.79788456080286535587989211986876f;
    return 0.5f * x * (1.0f + tanhf(s * x * (1.0f + c * x * x)));
}

template <int TYPE>
__device__ __forceinline__ void chain_stage(const ch_gemv & q, const hacts & a, uint32_t sbase, float * pw, int nr, int rsub, int RPP, bool active, int hh,
                                            int NW, int wir, int red, int lane) {
    for (int rr = rsub; rr - rsub < nr; rr += RPP) {
        float x = 0.0f;
        if (active && rr < nr) {
            hweights<TYPE> w;
            load_hweights<TYPE>(sbase + (uint32_t)rr * q.rb, hh, w);
            x = half_dot<TYPE, false>(w, a, hh, nullptr);
        }
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) if (o < red) x += __shfl_xor_sync(0xffffffffu, x, o);
        if ((lane & (red - 1)) == 0 && rr < nr) pw[rr * NW + wir] = x;
    }
}

__global__ void __launch_bounds__(288, 2) chain_kernel(const __grid_constant__ chain_params p) {
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    uint8_t * stages = smem_raw;
    uint8_t * act = smem_raw + (size_t)p.nstages * p.stage_bytes;                    // quantised column of the current phase; the fp32 row of a norm behind its first 6.4 KB
    float * wn = reinterpret_cast<float *>(act + p.act_bytes);                      // norm weights of the next norm
    float * partials = wn + p.wn_bytes / 4;                                          // [nstages][CH_PART]
    uint64_t * bars = reinterpret_cast<uint64_t *>(partials + (size_t)p.nstages * CH_PART);
    const uint32_t full0 = smem_u32(bars), empty0 = full0 + 8 * p.nstages, actbar = empty0 + 8 * p.nstages, wbar = actbar + 8;
    float * npart = reinterpret_cast<float *>(bars + 2 * p.nstages + 2);            // 32 floats
    float * rsd = npart + 32;                                                        // residual rows of this CTA (phases with epi 1)
    float * csn = rsd + p.rsd_rows;                                                  // (cos, sin) of this CTA's row pairs (the rope phase)

    if (tid == 0) {
        for (int s = 0; s < p.nstages; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 8); }
        mbar_init(actbar, 1);
        mbar_init(wbar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    pdl_trigger();

    if (warp == 0) {
        // ================= producer + epilogue =================
        const uint64_t pol = policy_evict_first();
        int ip = 0, is = 0, gsi = 0;                       // issue cursor: phase, stage in phase, global stage count
        ph_cta ic = phase_rows(p.ph[0]);
        while (ip < p.nph && ic.nst == 0) { ++ip; if (ip < p.nph) ic = phase_rows(p.ph[ip]); }
        auto issue_next = [&]() {                          // lane 0 only
            if (ip >= p.nph) return;
            const ch_gemv & q = p.ph[ip];
            const int slot = gsi % p.nstages;
            const int r0 = ic.row_lo + is * q.R, nr = min(q.R, ic.nrows - is * q.R);
            const uint32_t bytes = (uint32_t)nr * q.rb, bar = full0 + 8 * slot;
            mbar_expect_tx(bar, bytes);
            bulk_g2s_stream(smem_u32(stages + (size_t)slot * p.stage_bytes), q.W[ic.g] + (int64_t)r0 * q.rb, bytes, bar, pol);
            ++gsi;
            if (++is >= ic.nst) { is = 0; do { ++ip; if (ip < p.nph) ic = phase_rows(p.ph[ip]); } while (ip < p.nph && ic.nst == 0); }
        };
        if (lane == 0) for (int s = 0; s < p.nstages; ++s) issue_next();
        // cos / sin of the rope phase's row pairs: a function of the position (a graph input) only
        for (int ph = 0; ph < p.nph; ++ph) {
            const ch_gemv & q = p.ph[ph];
            if (q.epi != 2) continue;
            const ph_cta c = phase_rows(q);
            const int rmode = q.rmode[c.g];
            if (c.nst == 0 || (rmode != 1 && rmode != 2)) continue;
            const int hd = q.rhd[c.g], posv = q.rpos[0];
            for (int o2 = lane; 2 * o2 < c.nrows; o2 += 32) {
                const int d = (c.row_lo + 2 * o2) % hd;
                float cs = 1.0f, sn = 0.0f;
                if (d < q.rp.n_dims) rope_cs(q.rp, d >> 1, posv, q.rff, cs, sn);
                csn[2 * o2] = cs; csn[2 * o2 + 1] = sn;
            }
            __syncwarp();
        }
        pdl_wait();                                        // outputs may still be read by the previous kernel
        int gs = 0;
        for (int ph = 0; ph < p.nph; ++ph) {
            const ch_gemv & q = p.ph[ph];
            const ph_cta c = phase_rows(q);
            if (c.nst == 0) continue;                      // (a consumer thread feeds this phase's barrier for the CTA)
            const int NW = q.TPR > 32 ? q.TPR >> 5 : 1;
            float * Dc = q.D[c.g];
            const int rmode = q.epi == 2 ? q.rmode[c.g] : 0;
            for (int s = 0; s < c.nst; ++s, ++gs) {
                const int slot = gs % p.nstages;
                mbar_wait(empty0 + 8 * slot, (gs / p.nstages) & 1);
                if (s == 0 && lane == 0) ch_stamp(p, 32 + 2 * ph);
                if (s == 0 && q.epi == 1) {                // this CTA's rows of the residual (written before the barrier the consumers passed)
                    for (int o = lane; o < c.nrows; o += 32) rsd[o] = __ldcg(q.resid + c.row_lo + o);
                    __syncwarp();
                }
                const int r0 = c.row_lo + s * q.R, nr = min(q.R, c.nrows - s * q.R);
                const float * ps = partials + (size_t)slot * CH_PART;
                if (rmode == 0) {
                    for (int o = lane; o < nr; o += 32) {
                        float sum = 0.0f;
                        for (int wv = 0; wv < NW; ++wv) sum += ps[o * NW + wv];
                        Dc[r0 + o] = sum;
                        if (q.epi == 1) q.dsum[r0 + o] = sum + rsd[r0 - c.row_lo + o];
                    }
                } else {
                    char * cell = q.rslot[c.g] ? *q.rslot[c.g] : q.rcell[c.g];
                    for (int o0 = 0; o0 < nr; o0 += 32) {
                        const int o = o0 + lane, n = r0 + o;
                        const bool valid = o < nr;
                        float sum = 0.0f;
                        if (valid) for (int wv = 0; wv < NW; ++wv) sum += ps[o * NW + wv];
                        if (rmode == 3) {
                            if (valid) { Dc[n] = sum; store_as(cell + (size_t)n * q.rcesz[c.g], q.rctype[c.g], sum); }
                            continue;
                        }
                        const float other = __shfl_xor_sync(0xffffffffu, sum, 1);
                        if (valid) {
                            float y = sum;
                            if ((n % q.rhd[c.g]) < q.rp.n_dims) {
                                const int lo = (n - c.row_lo) & ~1;
                                float ya, yb;
                                rope_pair((n & 1) ? other : sum, (n & 1) ? sum : other, csn[lo], csn[lo + 1], ya, yb);
                                y = (n & 1) ? yb : ya;
                            }
                            q.rout[c.g][n] = y;
                            if (rmode == 2) store_as(cell + (size_t)n * q.rcesz[c.g], q.rctype[c.g], y);
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) {
                    if (s == c.nst - 1 && q.arrive_bar >= 0) grid_arrive(p, q.arrive_bar);
                    if (s == c.nst - 1) ch_stamp(p, 33 + 2 * ph);
                    issue_next();
                }
            }
        }
    } else {
        // ================= consumers =================
        const int ct = tid - 32, cw = warp - 1;
        hacts a;
        int gs = 0;
        uint32_t act_par = 0, w_par = 0;
        bool first_read = true;
        int norms_done = 0, w_issued = 0, w_waited = 0;
        if (p.norms[0].x) ++w_issued;
        if (ct == 0 && p.norms[0].x) { mbar_expect_tx(wbar, (uint32_t)p.norms[0].K * 4); bulk_g2s(smem_u32(wn), p.norms[0].w, (uint32_t)p.norms[0].K * 4, wbar); }
        for (int ph = 0; ph < p.nph; ++ph) {
            const ch_gemv & q = p.ph[ph];
            const ph_cta c = phase_rows(q);
            if (ct == 0 && ph == 0) ch_stamp(p, 0);
            if (q.silu_before) {
                // ---- act = unary(g), out = act * u, quantised column of `out`: 32-blocks dealt round-robin to the CTAs, 8 threads each ----
                if (ct == 0) { grid_wait(p, p.silu_wait); ch_stamp(p, 28); }
                asm volatile("bar.sync 1, 256;" ::: "memory");
                const int nblk = p.sn >> 5, sub = ct & 7;
                for (int base = 0; base < nblk; base += gridDim.x * 32) {      // (uniform trip count: quant_block_store shuffles over the full warp)
                    const int blk = base + blockIdx.x + gridDim.x * (ct >> 3);
                    const bool in = blk < nblk;
                    const int i = 8 * blk + sub;           // float4 index
                    float4 y = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (in) {
                        const float4 x = __ldcg(reinterpret_cast<const float4 *>(p.sg) + i);
                        const float4 uu = __ldcg(reinterpret_cast<const float4 *>(p.su) + i);
                        const float4 av = make_float4(chain_act(p.sop, x.x), chain_act(p.sop, x.y), chain_act(p.sop, x.z), chain_act(p.sop, x.w));
                        y = make_float4(av.x * uu.x, av.y * uu.y, av.z * uu.z, av.w * uu.w);
                        reinterpret_cast<float4 *>(p.sact)[i] = av;
                        reinterpret_cast<float4 *>(p.sout)[i] = y;
                    }
                    quant_block_store(y, in ? blk : 0, sub, in, p.sxq);
                }
                __threadfence();
                asm volatile("bar.sync 1, 256;" ::: "memory");
                if (ct == 0) { ch_stamp(p, 29); grid_arrive(p, p.silu_arrive); }
            }
            if (c.nst == 0 && !(q.pre == 1 && blockIdx.x == 0)) {      // nothing to multiply in this phase: only keep the barrier count right
                if (ct == 0 && q.arrive_bar >= 0) grid_arrive(p, q.arrive_bar);
                if (q.pre == 1) ++norms_done;
                continue;
            }
            if (q.pre != 2) {
                if (ct == 0) {
                    if (q.wait_bar >= 0) grid_wait(p, q.wait_bar);
                    else if (first_read) pdl_wait();       // the column comes from the previous kernel
                    ch_stamp(p, 1 + 4 * ph);
                    asm volatile("fence.proxy.async;" ::: "memory");
                    if (q.pre == 0) {
                        mbar_expect_tx(actbar, (uint32_t)q.BPR * ACT_ITEM);
                        bulk_g2s(smem_u32(act), q.xq, (uint32_t)q.BPR * ACT_ITEM, actbar);
                    } else {
                        const ch_norm & nm = p.norms[q.norm];
                        mbar_expect_tx(actbar, (uint32_t)nm.K * 4);
                        bulk_g2s(smem_u32(act) + (uint32_t)q.BPR * ACT_ITEM, nm.x, (uint32_t)nm.K * 4, actbar);
                    }
                }
                first_read = false;
                if (q.pre == 1) {
                    // RMS_NORM * w + q8 quantisation of the row: the arithmetic of add_rms_norm_mul_kernel<1> / mmvqr_kernel<.., NORM>
                    const ch_norm & nm = p.norms[q.norm];
                    const float * xs = reinterpret_cast<const float *>(act + (size_t)q.BPR * ACT_ITEM);
                    const int n4 = nm.K >> 2;
                    mbar_wait(wbar, w_par); w_par ^= 1; ++w_waited;
                    mbar_wait(actbar, act_par); act_par ^= 1;
                    if (ct == 0) ch_stamp(p, 44 + 4 * q.norm);
#pragma unroll 1
                    for (int k = 0; k < 4; ++k) {                                  // rolled: see mmvq_r.cuh (cold instruction cache)
                        const int i = ct + 256 * k;
                        if (i < n4) {
                            const float sq = warp_sum(sumsq4(reinterpret_cast<const float4 *>(xs)[i]));
                            if (lane == 0) npart[8 * k + cw] = sq;
                        }
                    }
                    asm volatile("bar.sync 1, 256;" ::: "memory");
                    if (ct == 0) ch_stamp(p, 45 + 4 * q.norm);
                    float tot = 0.0f;
                    if ((n4 >> 5) == 32) {                                     // K = 4096: loads batched, the same sequential order of additions
#pragma unroll
                        for (int w2 = 0; w2 < 16; ++w2) { const float2 t2 = reinterpret_cast<const float2 *>(npart)[w2]; tot += t2.x; tot += t2.y; }
                    } else {
                        for (int wi = 0; wi < (n4 >> 5); ++wi) tot += npart[wi];
                    }
                    if (ct == 0 && tot != -1.0f) ch_stamp(p, 52 + 4 * q.norm);
                    const float scale = rms_scale(tot, nm.K, nm.eps);
                    if (ct == 0 && scale != -1.0f) ch_stamp(p, 53 + 4 * q.norm);
                    const bool writer = blockIdx.x == 0;
                    {                                                          // eight consecutive elements per thread and pass: four lanes per 32-block
                        float4 y0[2], y1[2];
                        int jb[2];
                        bool in[2];
#pragma unroll
                        for (int k = 0; k < 2; ++k) {
                            const int t = ct + 256 * k, i = 2 * t;
                            in[k] = i < n4;
                            jb[k] = in[k] ? t >> 2 : 0;
                            y0[k] = make_float4(0.0f, 0.0f, 0.0f, 0.0f); y1[k] = y0[k];
                            if (in[k]) {
                                const float4 v0 = reinterpret_cast<const float4 *>(xs)[i], v1 = reinterpret_cast<const float4 *>(xs)[i + 1];
                                const float4 w0 = reinterpret_cast<const float4 *>(wn)[i], w1 = reinterpret_cast<const float4 *>(wn)[i + 1];
                                const float4 n0 = make_float4(v0.x * scale, v0.y * scale, v0.z * scale, v0.w * scale);
                                const float4 n1 = make_float4(v1.x * scale, v1.y * scale, v1.z * scale, v1.w * scale);
                                y0[k] = make_float4(n0.x * w0.x, n0.y * w0.y, n0.z * w0.z, n0.w * w0.w);
                                y1[k] = make_float4(n1.x * w1.x, n1.y * w1.y, n1.z * w1.z, n1.w * w1.w);
                                if (writer) {
                                    reinterpret_cast<float4 *>(nm.y)[i] = n0; reinterpret_cast<float4 *>(nm.y)[i + 1] = n1;
                                    reinterpret_cast<float4 *>(nm.out)[i] = y0[k]; reinterpret_cast<float4 *>(nm.out)[i + 1] = y1[k];
                                }
                            }
                        }
                        quant_block_store8<2>(y0, y1, jb, ct & 3, in, act);
                    }
                    if (ct == 0) ch_stamp(p, 46 + 4 * q.norm);
                    asm volatile("bar.sync 1, 256;" ::: "memory");
                    if (ct == 0) ch_stamp(p, 47 + 4 * q.norm);
                    ++norms_done;
                    if (norms_done < 2 && p.norms[norms_done].x) ++w_issued;
                    if (ct == 0 && norms_done < 2 && p.norms[norms_done].x) {      // the weights of the next norm take this one's place
                        asm volatile("fence.proxy.async;" ::: "memory");
                        mbar_expect_tx(wbar, (uint32_t)p.norms[norms_done].K * 4);
                        bulk_g2s(smem_u32(wn), p.norms[norms_done].w, (uint32_t)p.norms[norms_done].K * 4, wbar);
                    }
                } else {
                    mbar_wait(actbar, act_par); act_par ^= 1;
                }
            }
            if (c.nst == 0) { if (ct == 0 && q.arrive_bar >= 0) grid_arrive(p, q.arrive_bar); continue; }      // (CTA 0 of a norm phase always has rows; kept for safety)
            const int h = ct & (q.TPR - 1), rsub = ct >> q.lgTPR, RPP = 256 >> q.lgTPR;
            const bool active = h < 2 * q.BPR;
            const int b = h >> 1, hh = h & 1, wir = h >> 5, red = min(q.TPR, 32), NW = q.TPR > 32 ? q.TPR >> 5 : 1;
            if (q.pre != 2) {
                if (active) load_hacts(smem_u32(act) + (uint32_t)b * ACT_ITEM + hh * (4 * ACT_BLK), a);
                // the next phase's column may overwrite `act` only when every thread has its registers
                asm volatile("bar.sync 1, 256;" ::: "memory");
            }
            if (ct == 0) ch_stamp(p, 2 + 4 * ph);
            const uint32_t ioff = (uint32_t)b * (q.type == T_Q4_K ? (uint32_t)item_traits<T_Q4_K>::BYTES : (uint32_t)item_traits<T_Q6_K>::BYTES);
            for (int s = 0; s < c.nst; ++s, ++gs) {
                const int slot = gs % p.nstages;
                const int nr = min(q.R, c.nrows - s * q.R);
                mbar_wait(full0 + 8 * slot, (gs / p.nstages) & 1);
                if (s == 0 && ct == 0) ch_stamp(p, 3 + 4 * ph);
                const uint32_t sbase = smem_u32(stages + (size_t)slot * p.stage_bytes) + ioff;
                float * pw = partials + (size_t)slot * CH_PART;
                if (q.type == T_Q4_K) chain_stage<T_Q4_K>(q, a, sbase, pw, nr, rsub, RPP, active, hh, NW, wir, red, lane);
                else chain_stage<T_Q6_K>(q, a, sbase, pw, nr, rsub, RPP, active, hh, NW, wir, red, lane);
                __syncwarp();
                if (lane == 0) mbar_arrive(empty0 + 8 * slot);
            }
            if (ct == 0) ch_stamp(p, 4 + 4 * ph);
        }
        if (w_issued > w_waited) mbar_wait(wbar, w_par);      // a CTA that sat out the last norm: no bulk copy may be in flight into shared memory when it exits
    }
}

}  // namespace b200

using namespace b200;

namespace {

bool ch_debug() { static const bool v = [] { const char * e = getenv("B200_CHAIN_DEBUG"); return e && atoi(e) != 0; }(); return v; }
#define CH_FAIL(ret) do { if (ch_debug()) fprintf(stderr, "b200 chain: refused at %s:%d\n", __FILE__, __LINE__); return ret; } while (0)

int ch_sm_count() {
    static int n[16] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 16) dev = 0;
    if (!n[dev]) cudaDeviceGetAttribute(&n[dev], cudaDevAttrMultiProcessorCount, dev);
    return n[dev] > 0 ? n[dev] : 148;
}
int ch_env(const char * name, int dflt) { const char * e = getenv(name); return e ? atoi(e) : dflt; }
bool al16(const void * p) { return ((uintptr_t)p & 15) == 0; }
int round_up(int x, int m) { return (x + m - 1) / m * m; }

// geometry of one phase; returns false when the kernel cannot take it
bool ch_setup_gemv(const b200_chain_gemv & g, int grid, int stage_cap, ch_gemv & q, int * stage_bytes, int * rows_per_cta_out) {
    if (g.ng < 1 || g.ng > 3 || (g.type != T_Q4_K && g.type != T_Q6_K)) CH_FAIL(false);
    const int K = g.K;
    if (K <= 0 || K % 256) CH_FAIL(false);
    const int bpr = K / 256, hpr = 2 * bpr;
    if (hpr > 256) CH_FAIL(false);
    int tpr = 1, lg = 0;
    while (tpr < hpr) { tpr <<= 1; ++lg; }
    if (hpr * 10 < tpr * 6) CH_FAIL(false);
    const int rb = bpr * block_bytes(g.type);
    if (rb % 16) CH_FAIL(false);
    const int unit = g.epi == 2 ? 2 : 1;
    int ntot = 0;
    for (int i = 0; i < g.ng; ++i) {
        if (!al16(g.W[i]) || g.N[i] <= 0 || g.N[i] % unit || !g.dst[i]) CH_FAIL(false);
        ntot += g.N[i];
    }
    const int rpp = 256 / tpr, nw = tpr > 32 ? tpr / 32 : 1;
    q = ch_gemv();
    q.ng = g.ng; q.type = g.type; q.K = K; q.BPR = bpr; q.TPR = tpr; q.lgTPR = lg; q.rb = rb; q.epi = g.epi;
    int used = 0, rows_per_cta = 1;
    for (int i = 0; i < g.ng; ++i) {
        int share = i + 1 == g.ng ? std::max(1, grid - used) : std::max(1, (int)((int64_t)grid * g.N[i] / ntot));
        share = std::max(1, std::min(share, (g.N[i] / unit + rpp - 1) / rpp));
        if (used + share > grid) share = grid - used;
        if (share < 1) CH_FAIL(false);
        q.cta0[i] = used; used += share;
        q.W[i] = (const uint8_t *)g.W[i]; q.D[i] = g.dst[i];
        q.rows_q[i] = g.N[i] / unit / share; q.rows_rem[i] = g.N[i] / unit % share;
        rows_per_cta = std::max(rows_per_cta, unit * ((g.N[i] / unit + share - 1) / share));
    }
    for (int i = g.ng; i <= 3; ++i) q.cta0[i] = used;
    int R = std::max(rpp, (stage_cap / rb) / rpp * rpp);
    R = std::min(R, std::max(rpp, round_up((rows_per_cta + 3) / 4, rpp)));
    R = std::min(R, round_up(rows_per_cta, rpp));
    if (g.epi == 2 && (R & 1)) R = round_up(R, 2);
    if (R * nw > CH_PART || R * rb + 16 > stage_cap + 4096) CH_FAIL(false);
    q.R = R;
    *stage_bytes = std::max(*stage_bytes, round_up(R * rb + 16, 128));
    *rows_per_cta_out = rows_per_cta;
    if (g.epi == 1) { if (!g.resid || !g.sum || g.ng != 1) CH_FAIL(false); q.resid = g.resid; q.dsum = g.sum; }
    if (g.epi == 2) {
        const b200_gemv_rope * r = g.rope;
        if (!r || !r->pos || r->prm.mode != 0 || r->prm.n_dims % 2 || r->prm.n_dims <= 0) CH_FAIL(false);
        for (int i = 0; i < g.ng; ++i) {
            const int m = r->mode[i];
            if (m < 0 || m > 3) CH_FAIL(false);
            if ((m == 1 || m == 2) && (!r->out[i] || r->head_dim[i] <= 0 || r->head_dim[i] % 2 || r->prm.n_dims > r->head_dim[i] || g.N[i] % r->head_dim[i])) CH_FAIL(false);
            if ((m == 2 || m == 3) && (!(r->cell[i] || r->cell_slot[i]) || (r->cell_type[i] != T_F16 && r->cell_type[i] != T_F32 && r->cell_type[i] != T_BF16))) CH_FAIL(false);
            q.rmode[i] = m; q.rout[i] = r->out[i]; q.rcell[i] = (char *)r->cell[i]; q.rslot[i] = (char * const *)r->cell_slot[i];
            q.rctype[i] = r->cell_type[i]; q.rcesz[i] = r->cell_type[i] == T_F32 ? 4 : 2; q.rhd[i] = std::max(2, r->head_dim[i]);
        }
        q.rpos = r->pos; q.rff = r->ffac; q.rp = rope_dev_from(r->prm);
    }
    return true;
}

bool ch_norm_ok(const b200_chain_norm & n, int K) {
    return n.x && n.w && n.y && n.out && n.K == K && K <= 4096 && K % 256 == 0 && al16(n.x) && al16(n.w) && al16(n.y) && al16(n.out);
}

// grid: every CTA must be resident at once (they wait for each other)
int ch_grid(size_t smem) {
    static int per_sm[16] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev >= 16) dev = 0;
    if (!per_sm[dev]) {
        if (cudaFuncSetAttribute(chain_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448) != cudaSuccess) return 0;
        int n = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, chain_kernel, 288, 114688) != cudaSuccess) return 0;
        per_sm[dev] = std::max(0, std::min(2, n));
        if (!per_sm[dev]) per_sm[dev] = -1;
    }
    (void)smem;
    return per_sm[dev] > 0 ? per_sm[dev] * ch_sm_count() : 0;
}

int ch_build(const b200_chain_args * a, chain_params & p, dim3 * grid_out, size_t * smem_out) {
    if (!a || !a->barriers || !a->xq_attn || !a->xq_ffn || !al16(a->xq_attn) || !al16(a->xq_ffn)) CH_FAIL(-1);
    static const int stage_kb = ch_env("B200_CHAIN_STAGE_KB", 36), smem_kb = ch_env("B200_CHAIN_SMEM_KB", 112);
    const int grid = ch_grid(0);
    if (grid < 8) CH_FAIL(-1);
    const int stage_cap = stage_kb * 1024;
    p = chain_params();
    int stage_bytes = 0, rpc = 0, rsd_rows = 0, csn_rows = 0;
    const b200_chain_gemv * src[CH_MAXPH] = { &a->wo, &a->gate_up, &a->down, &a->t1, &a->t2 };
    const int nsrc = a->tail ? (a->t2.ng > 0 ? 5 : 4) : 3;
    if (a->wo.ng != 1 || a->wo.epi != 1 || a->gate_up.ng != 2 || a->gate_up.epi != 0 || a->down.ng != 1 || a->down.epi != 1) CH_FAIL(-1);
    if (a->gate_up.N[0] != a->gate_up.N[1] || a->gate_up.N[0] != a->down.K || a->gate_up.N[0] % 32) CH_FAIL(-1);
    if (a->gate_up.K != a->wo.N[0] || a->down.N[0] != a->wo.N[0]) CH_FAIL(-1);
    if (a->act_op != B200_UNARY_SILU && a->act_op != B200_UNARY_GELU) CH_FAIL(-1);
    if (!a->act || !a->act_out || !al16(a->act) || !al16(a->act_out) || !al16(a->gate_up.dst[0]) || !al16(a->gate_up.dst[1])) CH_FAIL(-1);
    if (!ch_norm_ok(a->norm1, a->gate_up.K)) CH_FAIL(-1);
    if (a->tail) {
        if (a->t1.ng < 1 || !ch_norm_ok(a->norm2, a->t1.K) || a->t1.K != a->down.N[0]) CH_FAIL(-1);
        if (a->t2.ng > 0 && a->t2.K != a->t1.K) CH_FAIL(-1);
    }
    int act_k = 0;
    for (int i = 0; i < nsrc; ++i) {
        int rows = 0;
        if (!ch_setup_gemv(*src[i], grid, stage_cap, p.ph[i], &stage_bytes, &rows)) CH_FAIL(-1);
        rpc = std::max(rpc, rows);
        if (src[i]->epi == 1) rsd_rows = std::max(rsd_rows, rows);
        if (src[i]->epi == 2) {      // one phase at most that rotates (its cos / sin table is computed once); V-only store phases need none
            bool rot = false;
            for (int g = 0; g < src[i]->ng; ++g) rot |= src[i]->rope->mode[g] == 1 || src[i]->rope->mode[g] == 2;
            if (rot) { if (csn_rows) CH_FAIL(-1); csn_rows = rows + 2; }
        }
        act_k = std::max(act_k, src[i]->K);
        if (i < 3 && p.ph[i].cta0[src[i]->ng] != grid) CH_FAIL(-1);      // every CTA takes part in the body of the chain (tiny matrices: not worth a chain)
    }
    p.nph = nsrc;
    // phase wiring: barrier 0 after wo, 1 after gate/up, 2 after the SILU step, 3 after down
    p.ph[0].pre = 0; p.ph[0].xq = (const uint8_t *)a->xq_attn; p.ph[0].wait_bar = -1; p.ph[0].arrive_bar = 0;
    p.ph[1].pre = 1; p.ph[1].norm = 0; p.ph[1].wait_bar = 0; p.ph[1].arrive_bar = 1;
    p.ph[2].pre = 0; p.ph[2].xq = (const uint8_t *)a->xq_ffn; p.ph[2].wait_bar = 2; p.ph[2].arrive_bar = 3; p.ph[2].silu_before = 1;
    p.silu_wait = 1; p.silu_arrive = 2;
    p.norms[0].x = a->norm1.x; p.norms[0].w = a->norm1.w; p.norms[0].y = a->norm1.y; p.norms[0].out = a->norm1.out; p.norms[0].eps = a->norm1.eps; p.norms[0].K = a->norm1.K;
    if (a->tail) {
        p.ph[3].pre = 1; p.ph[3].norm = 1; p.ph[3].wait_bar = 3; p.ph[3].arrive_bar = -1;
        if (nsrc == 5) { p.ph[4].pre = 2; p.ph[4].wait_bar = -1; p.ph[4].arrive_bar = -1; }
        p.norms[1].x = a->norm2.x; p.norms[1].w = a->norm2.w; p.norms[1].y = a->norm2.y; p.norms[1].out = a->norm2.out; p.norms[1].eps = a->norm2.eps; p.norms[1].K = a->norm2.K;
    }
    p.sg = a->gate_up.dst[0]; p.su = a->gate_up.dst[1]; p.sact = a->act; p.sout = a->act_out; p.sxq = (uint8_t *)a->xq_ffn; p.sn = a->gate_up.N[0]; p.sop = a->act_op;
    p.bars = a->barriers; p.error = a->error; p.timeout_ns = (unsigned long long)a->timeout_ms * 1000000ull;
    p.trace = (unsigned long long *)a->trace;
    // shared memory
    const int act_bytes = round_up(std::max((act_k / 256) * ACT_ITEM, (a->norm1.K / 256) * ACT_ITEM + a->norm1.K * 4), 128);
    const int wn_bytes = round_up(a->norm1.K * 4, 128);
    p.stage_bytes = stage_bytes; p.act_bytes = act_bytes; p.wn_bytes = wn_bytes; p.rsd_rows = round_up(rsd_rows + 1, 4); p.csn_rows = round_up(csn_rows + 2, 4);
    const size_t fixed = (size_t)act_bytes + wn_bytes + 16 * 2 + 128 + (size_t)(p.rsd_rows + p.csn_rows) * 4 + 64;
    const size_t per_stage = (size_t)stage_bytes + CH_PART * 4 + 16;
    const size_t budget = (size_t)smem_kb * 1024;
    if (budget < fixed + 2 * per_stage) CH_FAIL(-1);
    p.nstages = (int)std::min<size_t>(8, (budget - fixed) / per_stage);
    *smem_out = (size_t)p.nstages * per_stage + fixed;
    *grid_out = dim3(grid, 1, 1);
    return 0;
}

}  // namespace

extern "C" int b200_decode_chain_supported(const b200_chain_args * a) {
    chain_params p; dim3 g; size_t smem = 0;
    return ch_build(a, p, &g, &smem) == 0 ? 1 : 0;
}

extern "C" int b200_decode_chain(const b200_chain_args * a, void * stream) {
    chain_params p; dim3 grid; size_t smem = 0;
    if (ch_build(a, p, &grid, &smem) != 0) return -1;
    cudaLaunchConfig_t lc = {};
    lc.gridDim = grid; lc.blockDim = dim3(288, 1, 1); lc.dynamicSmemBytes = smem; lc.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = (a->pdl >= 0 ? a->pdl : pdl_default()) ? 1 : 0;
    B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, chain_kernel, p));
    ++g_launches;
    return 0;
}
