// fattn.cu -- GGML_OP_FLASH_ATTN_EXT over an f16 / q8_0 KV cache (vector form: decode and small batches).
//
// Replaces ggml_cuda_flash_attn_ext's vec kernels (ggml-cuda/fattn.cu:271-346, fattn-vec-f32.cuh,
// fattn-common.cuh:237-268,312-354,451-468) with one split-KV kernel + a combine kernel:
//   grid (q rows, heads, KV splits), 4 warps per CTA; a warp walks tiles of 32 KV positions.
//   scores: lane j owns position j of the tile and runs the whole dot product q . k_j (Q lives in
//           shared memory and is read as a broadcast; for q8_0 K, Q is quantised to int8 per 32
//           values once per CTA and the dot is dp4a, as the reference does on both CPU and GPU);
//   online softmax per tile with warp reductions (reference ggml-cpu/ops.cpp:7141-7209);
//   P.V:    lanes own D/32 output dims, V rows are read coalesced, p_j arrives by shuffle.
// K and V are strided views of the cache (nb1/nb2/nb3 honoured); GQA by index division.
#include "mmvq_units.cuh"
#include "../../include/b200_ops.h"

#include <algorithm>
#include <cmath>

namespace b200 {

struct fa_params {
    const char * q; const char * k; const char * v; const char * mask; char * dst;
    float * part;                          // [rows][heads][nsplit][2 + DV] when nsplit > 1
    int64_t nbq1, nbq2, nbq3, nbk1, nbk2, nbk3, nbv1, nbv2, nbv3, nbm1, nbd1;
    int N, n_head, ne3, n_kv, rk2, rk3, rv2, rv3;
    int nsplit, tiles_per_split;
    float scale, max_bias, softcap, m0, m1;
    unsigned n_head_log2;
    uint8_t * xq; uint32_t xq_acs;         // optional (DV = 128, ne3 = 1): the result also as q8 activations of the next GEMV, acs bytes per row
};

__device__ __forceinline__ uint32_t ldg_u16(const void * p) { return *reinterpret_cast<const unsigned short *>(p); }
__device__ __forceinline__ uint2 lds64(uint32_t saddr) {
    uint2 v; asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(saddr)); return v;
}

// dot product of one K row with Q (shared memory); D elements
template <int KT, int D>
__device__ __forceinline__ float k_dot(const char * krow, const float * qf, const int * qi, const float * qd) {
    float s = 0.0f;
    if constexpr (KT == T_F16) {
        const uint4 * kp = reinterpret_cast<const uint4 *>(krow);      // rows of the cache are 16-byte aligned
        uint4 kr[D / 8];                                               // the whole row in flight at once: one L2 round trip
#pragma unroll
        for (int i = 0; i < D / 8; ++i) kr[i] = __ldg(kp + i);
#pragma unroll
        for (int i = 0; i < D / 8; ++i) {
            const uint4 u = kr[i];
            const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&u.x)), b = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
            const float2 c = __half22float2(*reinterpret_cast<const __half2 *>(&u.z)), d = __half22float2(*reinterpret_cast<const __half2 *>(&u.w));
            const float4 q0 = *reinterpret_cast<const float4 *>(qf + 8 * i), q1 = *reinterpret_cast<const float4 *>(qf + 8 * i + 4);
            s = fmaf(a.x, q0.x, s); s = fmaf(a.y, q0.y, s); s = fmaf(b.x, q0.z, s); s = fmaf(b.y, q0.w, s);
            s = fmaf(c.x, q1.x, s); s = fmaf(c.y, q1.y, s); s = fmaf(d.x, q1.z, s); s = fmaf(d.y, q1.w, s);
        }
    } else {   // q8_0: blocks of {half d; int8 q[32]}, 2-byte aligned
#pragma unroll
        for (int b = 0; b < D / 32; ++b) {
            const char * blk = krow + 34 * b;
            const float dk = __half2float(__ushort_as_half((unsigned short)ldg_u16(blk)));
            int acc = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                const int w = (int)(ldg_u16(blk + 2 + 4 * i) | (ldg_u16(blk + 4 + 4 * i) << 16));
                acc = __dp4a(w, qi[8 * b + i], acc);
            }
            s = fmaf(dk * qd[b], (float)acc, s);
        }
    }
    return s;
}

// lanes own DV/32 consecutive output dims: acc[] += p * V[row][dims]
template <int VT, int DV>
__device__ __forceinline__ void v_accum(const char * vrow, int lane, float p, float (&acc)[DV / 32]) {
    constexpr int E = DV / 32;
    if constexpr (VT == T_F16) {
        const __half * vp = reinterpret_cast<const __half *>(vrow) + E * lane;
        if constexpr (E == 4) {
            const uint2 u = __ldg(reinterpret_cast<const uint2 *>(vp));
            const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&u.x)), b = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
            acc[0] = fmaf(p, a.x, acc[0]); acc[1] = fmaf(p, a.y, acc[1]); acc[2] = fmaf(p, b.x, acc[2]); acc[3] = fmaf(p, b.y, acc[3]);
        } else {
#pragma unroll
            for (int e = 0; e < E; ++e) acc[e] = fmaf(p, __half2float(vp[e]), acc[e]);
        }
    } else {
        const int d0 = E * lane;
        const char * blk = vrow + 34 * (d0 >> 5);
        const float dv = __half2float(__ushort_as_half((unsigned short)ldg_u16(blk)));
        const float pd = p * dv;
#pragma unroll
        for (int e = 0; e < E; e += 2) {
            const uint32_t w = ldg_u16(blk + 2 + (d0 & 31) + e);
            acc[e] = fmaf(pd, (float)(int)(int8_t)(w & 0xFF), acc[e]);
            acc[e + 1] = fmaf(pd, (float)(int)(int8_t)(w >> 8), acc[e + 1]);
        }
    }
}

// the same accumulation split into load and use, so that the loads of many KV positions can be in flight together
struct vfrag { uint32_t a, b; };
template <int VT, int DV>
__device__ __forceinline__ vfrag v_load(const char * vrow, int lane) {
    constexpr int E = DV / 32;
    vfrag f; f.a = 0; f.b = 0;
    if constexpr (VT == T_F16) {
        if constexpr (E == 4) { const uint2 u = __ldg(reinterpret_cast<const uint2 *>(vrow) + lane); f.a = u.x; f.b = u.y; }
        else f.a = __ldg(reinterpret_cast<const uint32_t *>(vrow) + lane);
    } else {
        const int d0 = E * lane;
        const char * blk = vrow + 34 * (d0 >> 5);
        f.b = ldg_u16(blk);
        f.a = ldg_u16(blk + 2 + (d0 & 31));
        if constexpr (E == 4) f.a |= ldg_u16(blk + 4 + (d0 & 31)) << 16;
    }
    return f;
}
template <int VT, int DV>
__device__ __forceinline__ void v_fma(const vfrag f, float p, float (&acc)[DV / 32]) {
    constexpr int E = DV / 32;
    static_assert(E == 2 || E == 4, "head sizes 64 and 128");
    if constexpr (VT == T_F16) {
        const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&f.a));
        acc[0] = fmaf(p, a.x, acc[0]); acc[1] = fmaf(p, a.y, acc[1]);
        if constexpr (E == 4) {
            const float2 b = __half22float2(*reinterpret_cast<const __half2 *>(&f.b));
            acc[2] = fmaf(p, b.x, acc[2]); acc[3] = fmaf(p, b.y, acc[3]);
        }
    } else {
        const float pd = p * __half2float(__ushort_as_half((unsigned short)f.b));
#pragma unroll
        for (int e = 0; e < E; ++e) acc[e] = fmaf(pd, (float)(int)(int8_t)((f.a >> (8 * e)) & 0xFF), acc[e]);
    }
}

// Decode at short context is pure latency: 8 warps so that up to 256 positions are one tile per warp, the mask value and
// the whole K row of a lane's position are requested together, and the V rows of 16 positions are in flight at a time
// (kernel timeline, profiles/r01_ktrace_tg_v4.txt: 11.5 us per layer with dependent loads -> see v5).
template <int KT, int VT, int DK, int DV>
__global__ void __launch_bounds__(256) fattn_vec_kernel(const fa_params p) {
    pdl_trigger(); pdl_wait();
    constexpr int EV = DV / 32, NW = 8;
    __shared__ __align__(16) float qf[DK];
    __shared__ int qi[DK / 4];
    __shared__ float qd[DK / 32];
    __shared__ float wm[NW], wl[NW];
    __shared__ float wacc[NW][DV];

    const int iq1 = blockIdx.x, iq2 = blockIdx.y % p.n_head, iq3 = blockIdx.y / p.n_head, split = blockIdx.z;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const float * qrow = reinterpret_cast<const float *>(p.q + iq1 * p.nbq1 + iq2 * p.nbq2 + iq3 * p.nbq3);
    for (int i = tid; i < DK; i += 32 * NW) qf[i] = qrow[i];
    __syncthreads();
    if constexpr (KT == T_Q8_0) {
        // quantise Q per 32 values (q8_0 arithmetic of the CPU path: d = amax/127, q = roundf(x/d))
        for (int b = warp; b < DK / 32; b += NW) {
            const float x = qf[32 * b + lane];
            const float amax = warp_max(fabsf(x));
            const float d = amax / 127.0f;
            const int q = amax == 0.0f ? 0 : (int)roundf(x / d);
            reinterpret_cast<int8_t *>(qi)[32 * b + lane] = (int8_t)q;
            if (lane == 0) qd[b] = d;
        }
        __syncthreads();
    }

    ktrace_event(2);
    const unsigned h = iq2;
    const float slope = p.max_bias > 0.0f ? (h < p.n_head_log2 ? powf(p.m0, (float)(h + 1)) : powf(p.m1, (float)(2 * (h - p.n_head_log2) + 1))) : 1.0f;
    const char * kbase = p.k + (iq2 / p.rk2) * p.nbk2 + (iq3 / p.rk3) * p.nbk3;
    const char * vbase = p.v + (iq2 / p.rv2) * p.nbv2 + (iq3 / p.rv3) * p.nbv3;
    const __half * mrow = p.mask ? reinterpret_cast<const __half *>(p.mask + iq1 * p.nbm1) : nullptr;

    float M = -INFINITY, L = 0.0f, acc[EV];
#pragma unroll
    for (int e = 0; e < EV; ++e) acc[e] = 0.0f;

    const int ntiles = (p.n_kv + 31) / 32;
    const int t0 = split * p.tiles_per_split, t1 = min(ntiles, t0 + p.tiles_per_split);
    for (int t = t0 + warp; t < t1; t += NW) {
        const int ic = 32 * t + lane;
        float s = -INFINITY;
        if (ic < p.n_kv) {
            // the K row is requested together with the mask value (a masked position costs a wasted row read, an unmasked
            // one saves a dependent round trip); its dot product is dropped by the select below
            const float mv = mrow ? slope * __half2float(mrow[ic]) : 0.0f;
            const float sd = k_dot<KT, DK>(kbase + (int64_t)ic * p.nbk1, qf, qi, qd) * p.scale;
            if (mv != -INFINITY) {
                s = sd;
                if (p.softcap != 0.0f) s = p.softcap * tanhf(s);
                s += mv;
            }
        }
        const float tmax = warp_max(s);
        ktrace_event(3);
        if (tmax == -INFINITY) continue;            // whole tile masked (warp-uniform)
        const float Mn = fmaxf(M, tmax);
        const float corr = expf(M - Mn);            // M = -inf -> 0
        const float pj = expf(s - Mn);              // s = -inf -> 0
        L = L * corr + warp_sum(pj);
#pragma unroll
        for (int e = 0; e < EV; ++e) acc[e] *= corr;
        M = Mn;
        ktrace_event(4);
        const int nvalid = min(32, p.n_kv - 32 * t);
        constexpr int VB = 16;                      // V rows in flight per lane
        for (int j0 = 0; j0 < nvalid; j0 += VB) {
            float pv[VB]; vfrag vf[VB];
#pragma unroll
            for (int jj = 0; jj < VB; ++jj) {
                pv[jj] = __shfl_sync(0xffffffffu, pj, (j0 + jj) & 31);
                if (j0 + jj >= nvalid) pv[jj] = 0.0f;
                vf[jj].a = 0; vf[jj].b = 0;
                if (pv[jj] != 0.0f) vf[jj] = v_load<VT, DV>(vbase + (int64_t)(32 * t + j0 + jj) * p.nbv1, lane);
            }
#pragma unroll
            for (int jj = 0; jj < VB; ++jj)
                if (pv[jj] != 0.0f) v_fma<VT, DV>(vf[jj], pv[jj], acc);
        }
    }

    // combine the warps
    ktrace_event(5);
    if (lane == 0) { wm[warp] = M; wl[warp] = L; }
#pragma unroll
    for (int e = 0; e < EV; ++e) wacc[warp][EV * lane + e] = acc[e];
    __syncthreads();
    float Mc = wm[0];
    for (int w = 1; w < NW; ++w) Mc = fmaxf(Mc, wm[w]);
    float Lc = 0.0f;
    for (int w = 0; w < NW; ++w) Lc += wm[w] == -INFINITY ? 0.0f : wl[w] * expf(wm[w] - Mc);
    const int64_t orow = ((int64_t)iq3 * p.N + iq1) * p.n_head + iq2;
    __shared__ __align__(16) float fin[DV];
    for (int d = tid; d < DV; d += 32 * NW) {
        float a = 0.0f;
        for (int w = 0; w < NW; ++w) a += wm[w] == -INFINITY ? 0.0f : wacc[w][d] * expf(wm[w] - Mc);
        if (p.nsplit == 1) { const float o = a / Lc; reinterpret_cast<float *>(p.dst + orow * p.nbd1)[d] = o; fin[d] = o; }
        else p.part[(orow * p.nsplit + split) * (2 + DV) + 2 + d] = a;
    }
    if constexpr (DV == 128) {
        if (p.xq && p.nsplit == 1) {             // head iq2 = blocks 4*iq2 .. 4*iq2+3 of row iq1 in the GEMV's activation layout
            __syncthreads();
            if (warp == 0) quant_block_store(*reinterpret_cast<const float4 *>(&fin[4 * lane]), 4 * iq2 + (lane >> 3), lane & 7, true, p.xq + (size_t)iq1 * p.xq_acs);
        }
    }
    if (p.nsplit > 1 && tid == 0) {
        p.part[(orow * p.nsplit + split) * (2 + DV)] = Mc;
        p.part[(orow * p.nsplit + split) * (2 + DV) + 1] = Lc;
    }
    ktrace_event(6);
}

// ---- staged form for an f16 cache (decode at short and medium context) ------------------------------------------------
// The kernel above lets every lane pull its own K row straight from global memory: 16 LDG.128 per lane that touch 32
// different lines each, 2 us of L1 tag throughput per CTA before the first dot product (kernel timeline, event marks of
// profiles/r01_ktrace_tg_v4.txt), and the V rows follow in dependent round trips.  Here a warp pulls the K and V rows of its tile with COALESCED 16-byte loads (a row per 16
// lanes), all 32 of them in flight as soon as the PDL wait is passed -- mask, Q and the whole tile travel in the same
// round trip -- parks them in padded shared-memory rows, and the arithmetic then runs out of shared memory in exactly the
// order of fattn_vec_kernel (results are bit-identical to it).  (Per-lane 1-D bulk copies were tried first: 512 TMA
// operations per CTA take 2 us to issue.)
template <int D>
__global__ void __launch_bounds__(256) fattn_vec_stage_kernel(const fa_params p) {
    constexpr int EV = D / 32, NW = 8;
    constexpr int KROW = 2 * D + 16;              // padded: lanes reading consecutive rows with LDS.128 hit distinct banks
    constexpr int VROW = 2 * D;
    constexpr int WARP_BYTES = 32 * (KROW + VROW);
    extern __shared__ __align__(128) uint8_t fa_smem[];
    __shared__ __align__(16) float qf[D];
    __shared__ float wm[NW], wl[NW], wsc[NW];
    __shared__ float wacc[NW][D];
    __shared__ __align__(16) float fin[D];

    const int iq1 = blockIdx.x, iq2 = blockIdx.y % p.n_head, iq3 = blockIdx.y / p.n_head, split = blockIdx.z;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    pdl_trigger(); pdl_wait();

    const char * kbase = p.k + (iq2 / p.rk2) * p.nbk2 + (iq3 / p.rk3) * p.nbk3;
    const char * vbase = p.v + (iq2 / p.rv2) * p.nbv2 + (iq3 / p.rv3) * p.nbv3;
    uint8_t * kt = fa_smem + (size_t)warp * WARP_BYTES, * vt = kt + 32 * KROW;
    const int ntiles = (p.n_kv + 31) / 32;
    const int t0 = split * p.tiles_per_split, t1 = min(ntiles, t0 + p.tiles_per_split);
    // one tile = 32 positions: coalesced 16-byte loads (a row of 2*D bytes per 2*D/16 lanes), all in flight at once, then
    // parked in shared memory where lane j finds row j
    constexpr int LPR = 2 * D / 16, RPI = 32 / LPR, NI = 32 / RPI;     // lanes per row, rows per instruction, instructions
    uint4 kreg[NI], vreg[NI];
    auto fetch = [&](int t) {
        const int nvalid = min(32, p.n_kv - 32 * t);
#pragma unroll
        for (int i = 0; i < NI; ++i) {
            const int r = RPI * i + lane / LPR;
            kreg[i] = make_uint4(0, 0, 0, 0); vreg[i] = make_uint4(0, 0, 0, 0);
            if (r < nvalid) {
                kreg[i] = __ldg(reinterpret_cast<const uint4 *>(kbase + (int64_t)(32 * t + r) * p.nbk1) + lane % LPR);
                vreg[i] = __ldg(reinterpret_cast<const uint4 *>(vbase + (int64_t)(32 * t + r) * p.nbv1) + lane % LPR);
            }
        }
    };
    auto park = [&]() {
#pragma unroll
        for (int i = 0; i < NI; ++i) {
            const int r = RPI * i + lane / LPR;
            *reinterpret_cast<uint4 *>(kt + r * KROW + 16 * (lane % LPR)) = kreg[i];
            *reinterpret_cast<uint4 *>(vt + r * VROW + 16 * (lane % LPR)) = vreg[i];
        }
        __syncwarp();
    };
    if (t0 + warp < t1) fetch(t0 + warp);

    const float * qrow = reinterpret_cast<const float *>(p.q + iq1 * p.nbq1 + iq2 * p.nbq2 + iq3 * p.nbq3);
    for (int i = tid; i < D; i += 32 * NW) qf[i] = qrow[i];
    const unsigned h = iq2;
    const float slope = p.max_bias > 0.0f ? (h < p.n_head_log2 ? powf(p.m0, (float)(h + 1)) : powf(p.m1, (float)(2 * (h - p.n_head_log2) + 1))) : 1.0f;
    const __half * mrow = p.mask ? reinterpret_cast<const __half *>(p.mask + iq1 * p.nbm1) : nullptr;
    __syncthreads();
    ktrace_event(2);

    float M = -INFINITY, L = 0.0f, acc[EV];
#pragma unroll
    for (int e = 0; e < EV; ++e) acc[e] = 0.0f;
    for (int t = t0 + warp; t < t1; t += NW) {
        const int ic = 32 * t + lane;
        const float mv = (ic < p.n_kv && mrow) ? slope * __half2float(mrow[ic]) : 0.0f;
        if (t != t0 + warp) fetch(t);
        park();
        float s = -INFINITY;
        if (ic < p.n_kv && mv != -INFINITY) {
            const uint32_t kr = smem_u32(kt + lane * KROW);
            float sd = 0.0f;
#pragma unroll
            for (int i = 0; i < D / 8; ++i) {
                const int4 u = lds128(kr + 16 * i);
                const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&u.x)), b = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
                const float2 c = __half22float2(*reinterpret_cast<const __half2 *>(&u.z)), d = __half22float2(*reinterpret_cast<const __half2 *>(&u.w));
                const float4 q0 = *reinterpret_cast<const float4 *>(qf + 8 * i), q1 = *reinterpret_cast<const float4 *>(qf + 8 * i + 4);
                sd = fmaf(a.x, q0.x, sd); sd = fmaf(a.y, q0.y, sd); sd = fmaf(b.x, q0.z, sd); sd = fmaf(b.y, q0.w, sd);
                sd = fmaf(c.x, q1.x, sd); sd = fmaf(c.y, q1.y, sd); sd = fmaf(d.x, q1.z, sd); sd = fmaf(d.y, q1.w, sd);
            }
            s = sd * p.scale;
            if (p.softcap != 0.0f) s = p.softcap * tanhf(s);
            s += mv;
        }
        const float tmax = warp_max(s);
        ktrace_event(3);
        if (tmax != -INFINITY) {                    // else: whole tile masked (warp-uniform)
            const float Mn = fmaxf(M, tmax);
            const float corr = expf(M - Mn);            // M = -inf -> 0
            const float pj = expf(s - Mn);              // s = -inf -> 0
            L = L * corr + warp_sum(pj);
#pragma unroll
            for (int e = 0; e < EV; ++e) acc[e] *= corr;
            M = Mn;
            const int nvalid = min(32, p.n_kv - 32 * t);
            const uint32_t vr = smem_u32(vt) + 2 * EV * lane;
#pragma unroll 8
            for (int j = 0; j < nvalid; ++j) {
                const float pv = __shfl_sync(0xffffffffu, pj, j);
                vfrag f; f.b = 0;
                if constexpr (EV == 4) { f.a = lds32(vr + j * VROW); f.b = lds32(vr + j * VROW + 4); } else f.a = lds32(vr + j * VROW);
                if (pv != 0.0f) v_fma<T_F16, D>(f, pv, acc);
            }
        }
        __syncwarp();
    }
    ktrace_event(5);

    // combine the warps
    if (lane == 0) { wm[warp] = M; wl[warp] = L; }
#pragma unroll
    for (int e = 0; e < EV; ++e) wacc[warp][EV * lane + e] = acc[e];
    __syncthreads();
    float Mc = wm[0];
    for (int w = 1; w < NW; ++w) Mc = fmaxf(Mc, wm[w]);
    if (tid < NW) wsc[tid] = wm[tid] == -INFINITY ? 0.0f : expf(wm[tid] - Mc);
    __syncthreads();
    float Lc = 0.0f;
    for (int w = 0; w < NW; ++w) Lc += wm[w] == -INFINITY ? 0.0f : wl[w] * wsc[w];
    const int64_t orow = ((int64_t)iq3 * p.N + iq1) * p.n_head + iq2;
    for (int d = tid; d < D; d += 32 * NW) {
        float a = 0.0f;
        for (int w = 0; w < NW; ++w) a += wm[w] == -INFINITY ? 0.0f : wacc[w][d] * wsc[w];
        if (p.nsplit == 1) { const float o = a / Lc; reinterpret_cast<float *>(p.dst + orow * p.nbd1)[d] = o; fin[d] = o; }
        else p.part[(orow * p.nsplit + split) * (2 + D) + 2 + d] = a;
    }
    if constexpr (D == 128) {
        if (p.xq && p.nsplit == 1) {             // head iq2 = blocks 4*iq2 .. 4*iq2+3 of row iq1 in the GEMV's activation layout
            __syncthreads();
            if (warp == 0) quant_block_store(*reinterpret_cast<const float4 *>(&fin[4 * lane]), 4 * iq2 + (lane >> 3), lane & 7, true, p.xq + (size_t)iq1 * p.xq_acs);
        }
    }
    if (p.nsplit > 1 && tid == 0) {
        p.part[(orow * p.nsplit + split) * (2 + D)] = Mc;
        p.part[(orow * p.nsplit + split) * (2 + D) + 1] = Lc;
    }
    ktrace_event(6);
}

// ---- grouped form: G query vectors that share one KV head (GQA heads x the rows of a small batch) per CTA ----
// At long context the kernel above reads the cache once per (row, head): rk2 x N times more L2 traffic than the cache
// holds.  Here lane j still owns KV position j of a 32-position tile, but its K row is loaded once and dotted with all G
// queries (Q in shared memory, broadcast reads), and every V row chunk is loaded once and accumulated into G outputs.
// Query g of CTA x: flat index x * G + g over (row, head-in-group), row-major.
// STAGE (f16 cache, 16-byte aligned rows): the K rows of a warp's tile reach shared memory through cp.async, a row per 2D/16 lanes
// (whole 128-byte lines per request instead of 32 scattered 16-byte pieces per load instruction), double buffered per warp so the
// next tile is in flight while this one is multiplied; lane j then reads row j from padded shared-memory rows (conflict-free
// LDS.128).  Same arithmetic in the same order as the direct form: results are bit-identical to it.
// ncu (profiles/r02_fattn_gqa_32k_ncu.md): with K alone staged and V read straight from global memory the kernel ran at 1.67 TB/s
// with 8 warps per SM, 28 % issue slots, stalled on the scoreboard: too few bytes in flight.  So V goes through the same
// cp.async pipeline, a CTA is 6 warps that own a whole SM (one K + V tile of 16 KB in flight per warp while the previous one is
// multiplied), and the per-warp combine buffers live in the staging memory once the tiles are done.
template <int KT, int VT, int D, int G, bool STAGE = false>
__global__ void __launch_bounds__(STAGE ? 192 : 128) fattn_vec_gqa_kernel(const fa_params p) {
    pdl_trigger(); pdl_wait();
    constexpr int EV = D / 32, NW = STAGE ? 6 : 4, NT = 32 * NW;
    constexpr int KROW = 2 * D + 16, VROW = 2 * D, TILEB = 32 * (KROW + VROW);
    extern __shared__ __align__(128) uint8_t gqa_smem[];
    __shared__ __align__(16) float qf[G][D];
    __shared__ int qi[G][D / 4];
    __shared__ float qd[G][D / 32];
    __shared__ float wm[NW][G], wl[NW][G];
    __shared__ float wacc_s[STAGE ? 1 : NW][STAGE ? 1 : G][STAGE ? 1 : D];
    float (*wacc)[G][D] = STAGE ? reinterpret_cast<float (*)[G][D]>(gqa_smem) : reinterpret_cast<float (*)[G][D]>(&wacc_s[0][0][0]);

    const int ikv = blockIdx.y % (p.n_head / p.rk2), iq3 = blockIdx.y / (p.n_head / p.rk2), split = blockIdx.z;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int ngq = p.N * p.rk2;                                  // queries that share this KV head
    const int g0 = blockIdx.x * G;
    for (int i = tid; i < G * D; i += NT) {
        const int g = i / D, d = i % D, gq = g0 + g;
        float x = 0.0f;
        if (gq < ngq) x = reinterpret_cast<const float *>(p.q + (int64_t)(gq / p.rk2) * p.nbq1 + (int64_t)(ikv * p.rk2 + gq % p.rk2) * p.nbq2 + (int64_t)iq3 * p.nbq3)[d];
        qf[g][d] = x;
    }
    __syncthreads();
    if constexpr (KT == T_Q8_0) {
        for (int b = warp; b < G * (D / 32); b += NW) {
            const int g = b / (D / 32), bb = b % (D / 32);
            const float x = qf[g][32 * bb + lane];
            const float amax = warp_max(fabsf(x));
            const float d = amax / 127.0f;
            const int q = amax == 0.0f ? 0 : (int)roundf(x / d);
            reinterpret_cast<int8_t *>(qi[g])[32 * bb + lane] = (int8_t)q;
            if (lane == 0) qd[g][bb] = d;
        }
        __syncthreads();
    }
    float slope[G];
    const __half * mrow[G];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        const int gq = min(g0 + g, ngq - 1);
        const unsigned h = (unsigned)(ikv * p.rk2 + gq % p.rk2);
        slope[g] = p.max_bias > 0.0f ? (h < p.n_head_log2 ? powf(p.m0, (float)(h + 1)) : powf(p.m1, (float)(2 * (h - p.n_head_log2) + 1))) : 1.0f;
        mrow[g] = p.mask ? reinterpret_cast<const __half *>(p.mask + (int64_t)(gq / p.rk2) * p.nbm1) : nullptr;
    }
    const char * kbase = p.k + (int64_t)ikv * p.nbk2 + (int64_t)(iq3 / p.rk3) * p.nbk3;
    const char * vbase = p.v + (int64_t)ikv * p.nbv2 + (int64_t)(iq3 / p.rv3) * p.nbv3;

    float M[G], L[G], acc[G][EV];
#pragma unroll
    for (int g = 0; g < G; ++g) {
        M[g] = -INFINITY; L[g] = 0.0f;
#pragma unroll
        for (int e = 0; e < EV; ++e) acc[g][e] = 0.0f;
    }
    const int ntiles = (p.n_kv + 31) / 32;
    const int t0 = split * p.tiles_per_split, t1 = min(ntiles, t0 + p.tiles_per_split);
    constexpr int LPR = 2 * D / 16, RPI = 32 / LPR, NI = 32 / RPI;      // lanes per row, rows per instruction, instructions per tile
    uint8_t * kbuf = gqa_smem + (size_t)warp * 2 * TILEB;      // per warp: two tiles of (32 padded K rows, 32 V rows)
    auto prefetch = [&](int t, int b) {
        if constexpr (STAGE) {
            const int nvalid = min(32, p.n_kv - 32 * t);
#pragma unroll
            for (int i = 0; i < NI; ++i) {
                const int r = RPI * i + lane / LPR;
                if (r < nvalid) {
                    const uint32_t dk = smem_u32(kbuf + (size_t)b * TILEB + r * KROW + 16 * (lane % LPR));
                    const uint32_t dv = smem_u32(kbuf + (size_t)b * TILEB + 32 * KROW + r * VROW + 16 * (lane % LPR));
                    const char * sk = kbase + (int64_t)(32 * t + r) * p.nbk1 + 16 * (lane % LPR);
                    const char * sv = vbase + (int64_t)(32 * t + r) * p.nbv1 + 16 * (lane % LPR);
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dk), "l"(sk) : "memory");
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(dv), "l"(sv) : "memory");
                }
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
    };
    int kb = 0;
    if (t0 + warp < t1) prefetch(t0 + warp, 0);
    for (int t = t0 + warp; t < t1; t += NW, kb ^= 1) {
        const int ic = 32 * t + lane;
        float s[G];
#pragma unroll
        for (int g = 0; g < G; ++g) s[g] = 0.0f;
        const bool in = ic < p.n_kv;
        if constexpr (STAGE) {
            __syncwarp();                               // every lane is through with the tile whose buffer the next copy overwrites
            if (t + NW < t1) { prefetch(t + NW, kb ^ 1); asm volatile("cp.async.wait_group 1;" ::: "memory"); }
            else asm volatile("cp.async.wait_group 0;" ::: "memory");
            __syncwarp();
            if (in) {
                const uint32_t kr = smem_u32(kbuf + (size_t)kb * TILEB + lane * KROW);
#pragma unroll 4
                for (int i = 0; i < D / 8; ++i) {
                    const int4 u = lds128(kr + 16 * i);
                    const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&u.x)), b = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
                    const float2 c = __half22float2(*reinterpret_cast<const __half2 *>(&u.z)), d = __half22float2(*reinterpret_cast<const __half2 *>(&u.w));
#pragma unroll
                    for (int g = 0; g < G; ++g) {
                        const float4 q0 = *reinterpret_cast<const float4 *>(&qf[g][8 * i]), q1 = *reinterpret_cast<const float4 *>(&qf[g][8 * i + 4]);
                        float x = s[g];
                        x = fmaf(a.x, q0.x, x); x = fmaf(a.y, q0.y, x); x = fmaf(b.x, q0.z, x); x = fmaf(b.y, q0.w, x);
                        x = fmaf(c.x, q1.x, x); x = fmaf(c.y, q1.y, x); x = fmaf(d.x, q1.z, x); x = fmaf(d.y, q1.w, x);
                        s[g] = x;
                    }
                }
            }
        } else
        if (in) {
            const char * krow = kbase + (int64_t)ic * p.nbk1;
            if constexpr (KT == T_F16) {
                const uint4 * kp = reinterpret_cast<const uint4 *>(krow);
#pragma unroll 4
                for (int i = 0; i < D / 8; ++i) {
                    const uint4 u = __ldg(kp + i);
                    const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&u.x)), b = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
                    const float2 c = __half22float2(*reinterpret_cast<const __half2 *>(&u.z)), d = __half22float2(*reinterpret_cast<const __half2 *>(&u.w));
#pragma unroll
                    for (int g = 0; g < G; ++g) {
                        const float4 q0 = *reinterpret_cast<const float4 *>(&qf[g][8 * i]), q1 = *reinterpret_cast<const float4 *>(&qf[g][8 * i + 4]);
                        float x = s[g];
                        x = fmaf(a.x, q0.x, x); x = fmaf(a.y, q0.y, x); x = fmaf(b.x, q0.z, x); x = fmaf(b.y, q0.w, x);
                        x = fmaf(c.x, q1.x, x); x = fmaf(c.y, q1.y, x); x = fmaf(d.x, q1.z, x); x = fmaf(d.y, q1.w, x);
                        s[g] = x;
                    }
                }
            } else {
#pragma unroll
                for (int b = 0; b < D / 32; ++b) {
                    const char * blk = krow + 34 * b;
                    const float dk = __half2float(__ushort_as_half((unsigned short)ldg_u16(blk)));
                    int w[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) w[i] = (int)(ldg_u16(blk + 2 + 4 * i) | (ldg_u16(blk + 4 + 4 * i) << 16));
#pragma unroll
                    for (int g = 0; g < G; ++g) {
                        int a = 0;
#pragma unroll
                        for (int i = 0; i < 8; ++i) a = __dp4a(w[i], qi[g][8 * b + i], a);
                        s[g] = fmaf(dk * qd[g][b], (float)a, s[g]);
                    }
                }
            }
        }
        float pj[G];
        bool any = false;
#pragma unroll
        for (int g = 0; g < G; ++g) {
            float x = -INFINITY;
            if (in) {
                const float mv = mrow[g] ? slope[g] * __half2float(mrow[g][ic]) : 0.0f;
                if (mv != -INFINITY) {
                    x = s[g] * p.scale;
                    if (p.softcap != 0.0f) x = p.softcap * tanhf(x);
                    x += mv;
                }
            }
            const float tmax = warp_max(x);
            if (tmax == -INFINITY) { pj[g] = 0.0f; continue; }          // this query sees nothing in the tile (warp-uniform)
            any = true;
            const float Mn = fmaxf(M[g], tmax);
            const float corr = expf(M[g] - Mn);
            pj[g] = expf(x - Mn);
            L[g] = L[g] * corr + warp_sum(pj[g]);
#pragma unroll
            for (int e = 0; e < EV; ++e) acc[g][e] *= corr;
            M[g] = Mn;
        }
        if (!any) continue;
        const int nvalid = min(32, p.n_kv - 32 * t);
        auto vstep = [&](int j) {
            float vv[EV];
            const char * vrow = vbase + (int64_t)(32 * t + j) * p.nbv1;
            if constexpr (STAGE) {
                const uint32_t vr = smem_u32(kbuf + (size_t)kb * TILEB + 32 * KROW + j * VROW) + 2 * EV * lane;
                if constexpr (EV == 4) {
                    const uint2 u = lds64(vr);
                    const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&u.x)), b = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
                    vv[0] = a.x; vv[1] = a.y; vv[2] = b.x; vv[3] = b.y;
                } else {
                    const uint32_t u = lds32(vr);
                    const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&u));
                    vv[0] = a.x; vv[1] = a.y;
                }
            } else if constexpr (VT == T_F16) {
                const __half * vp = reinterpret_cast<const __half *>(vrow) + EV * lane;
                if constexpr (EV == 4) {
                    const uint2 u = __ldg(reinterpret_cast<const uint2 *>(vp));
                    const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&u.x)), b = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
                    vv[0] = a.x; vv[1] = a.y; vv[2] = b.x; vv[3] = b.y;
                } else {
#pragma unroll
                    for (int e = 0; e < EV; ++e) vv[e] = __half2float(vp[e]);
                }
            } else {
                const int d0 = EV * lane;
                const char * blk = vrow + 34 * (d0 >> 5);
                const float dv = __half2float(__ushort_as_half((unsigned short)ldg_u16(blk)));
#pragma unroll
                for (int e = 0; e < EV; e += 2) {
                    const uint32_t w = ldg_u16(blk + 2 + (d0 & 31) + e);
                    vv[e] = dv * (float)(int)(int8_t)(w & 0xFF); vv[e + 1] = dv * (float)(int)(int8_t)(w >> 8);
                }
            }
#pragma unroll
            for (int g = 0; g < G; ++g) {
                const float pv = __shfl_sync(0xffffffffu, pj[g], j);
#pragma unroll
                for (int e = 0; e < EV; ++e) acc[g][e] = fmaf(pv, vv[e], acc[g][e]);
            }
        };
        if (nvalid == 32) {
#pragma unroll 8
            for (int j = 0; j < 32; ++j) vstep(j);
        } else {
            for (int j = 0; j < nvalid; ++j) vstep(j);
        }
    }

    // combine the warps, one query at a time
    if constexpr (STAGE) __syncthreads();             // the combine buffers take the place of the staged tiles
#pragma unroll
    for (int g = 0; g < G; ++g) {
        if (lane == 0) { wm[warp][g] = M[g]; wl[warp][g] = L[g]; }
#pragma unroll
        for (int e = 0; e < EV; ++e) wacc[warp][g][EV * lane + e] = acc[g][e];
    }
    __syncthreads();
    for (int i = tid; i < G * D; i += NT) {
        const int g = i / D, d = i % D, gq = g0 + g;
        if (gq >= ngq) continue;
        float Mc = wm[0][g];
        for (int w = 1; w < NW; ++w) Mc = fmaxf(Mc, wm[w][g]);
        float Lc = 0.0f, a = 0.0f;
        for (int w = 0; w < NW; ++w) {
            if (wm[w][g] == -INFINITY) continue;
            const float f = expf(wm[w][g] - Mc);
            Lc += wl[w][g] * f; a += wacc[w][g][d] * f;
        }
        const int64_t orow = ((int64_t)iq3 * p.N + gq / p.rk2) * p.n_head + ikv * p.rk2 + gq % p.rk2;
        if (p.nsplit == 1) reinterpret_cast<float *>(p.dst + orow * p.nbd1)[d] = a / Lc;
        else {
            float * pr = p.part + (orow * p.nsplit + split) * (2 + D);
            pr[2 + d] = a;
            if (d == 0) { pr[0] = Mc; pr[1] = Lc; }
        }
    }
}

// merge the split-KV partials: one warp per output row
template <int DV>
__global__ void __launch_bounds__(128) fattn_combine_kernel(const float * __restrict__ part, char * __restrict__ dst, int64_t nbd1, int64_t nrows, int nsplit,
                                                            uint8_t * __restrict__ xq, uint32_t xq_acs, int n_head) {
    pdl_trigger(); pdl_wait();
    const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (r >= nrows) return;
    const float * pr = part + r * nsplit * (2 + DV);
    float Mc = -INFINITY;
    for (int s = 0; s < nsplit; ++s) Mc = fmaxf(Mc, pr[s * (2 + DV)]);
    float Lc = 0.0f, a[DV / 32];
#pragma unroll
    for (int e = 0; e < DV / 32; ++e) a[e] = 0.0f;
    for (int s = 0; s < nsplit; ++s) {
        const float m = pr[s * (2 + DV)];
        if (m == -INFINITY) continue;
        const float w = expf(m - Mc);
        Lc += pr[s * (2 + DV) + 1] * w;
#pragma unroll
        for (int e = 0; e < DV / 32; ++e) a[e] += pr[s * (2 + DV) + 2 + (DV / 32) * lane + e] * w;
    }
    float * out = reinterpret_cast<float *>(dst + r * nbd1);
#pragma unroll
    for (int e = 0; e < DV / 32; ++e) { a[e] = a[e] / Lc; out[(DV / 32) * lane + e] = a[e]; }
    if constexpr (DV == 128) {
        if (xq) {                                // r = row * n_head + head (ne3 = 1); the warp is uniform in r
            const int head = (int)(r % n_head);
            quant_block_store(make_float4(a[0], a[1], a[2], a[3]), 4 * head + (lane >> 3), lane & 7, true, xq + (size_t)(r / n_head) * xq_acs);
        }
    }
}

static int fa_splits(int64_t rows_heads, int ntiles) {
    // enough CTAs for ~2 per SM, at least 4 tiles (one per warp) per CTA
    // short contexts: one CTA per (row, head) walks all tiles (no combine launch); long: split for occupancy
    if (ntiles <= 32) return 1;
    int ns = (int)std::min<int64_t>(std::max<int64_t>(1, (2 * 148 + rows_heads - 1) / rows_heads), std::max(1, ntiles / 8));
    return std::max(1, std::min(ns, 64));
}

template <int KT, int VT, int D>
static int fa_launch(const fa_params & p, dim3 grid, cudaStream_t st) {
    cudaLaunchConfig_t lc = {};
    lc.gridDim = grid; lc.blockDim = dim3(256); lc.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = pdl_default() ? 1 : 0;
    B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, fattn_vec_kernel<KT, VT, D, D>, p));
    ++g_launches;
    if (p.nsplit > 1) {
        const int64_t nrows = (int64_t)p.N * p.n_head * p.ne3;
        lc.blockDim = dim3(128);
        lc.gridDim = dim3((unsigned)((nrows * 32 + 127) / 128));
        B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, fattn_combine_kernel<D>, (const float *)p.part, p.dst, p.nbd1, nrows, p.nsplit, p.xq, p.xq_acs, p.n_head));
        ++g_launches;
    }
    return 0;
}

template <int D>
static int fa_launch_stage(const fa_params & p, dim3 grid, cudaStream_t st) {
    constexpr size_t smem = 8 * 32 * (size_t)(2 * D + 16 + 2 * D);
    static int attr_set[16] = {0};   // per device
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 16 && !attr_set[dev]) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(fattn_vec_stage_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set[dev] = 1;
    }
    cudaLaunchConfig_t lc = {};
    lc.gridDim = grid; lc.blockDim = dim3(256); lc.dynamicSmemBytes = smem; lc.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = pdl_default() ? 1 : 0;
    B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, fattn_vec_stage_kernel<D>, p));
    ++g_launches;
    if (p.nsplit > 1) {
        const int64_t nrows = (int64_t)p.N * p.n_head * p.ne3;
        lc.blockDim = dim3(128); lc.dynamicSmemBytes = 0;
        lc.gridDim = dim3((unsigned)((nrows * 32 + 127) / 128));
        B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, fattn_combine_kernel<D>, (const float *)p.part, p.dst, p.nbd1, nrows, p.nsplit, p.xq, p.xq_acs, p.n_head));
        ++g_launches;
    }
    return 0;
}
static int fa_stage_on() { static const int v = [] { const char * e = getenv("B200_FATTN_STAGE"); return e ? atoi(e) : 1; }();  return v; }

// off by default: measured SLOWER than the direct form (32k context, 32/8 heads: 0.132 vs 0.119 ms per call).  The ncu capture
// (profiles/r02_fattn_gqa_32k_ncu.md) shows why staging cannot help: DRAM traffic already equals the cache bytes and the kernel
// issues 2900 warp instructions per 16 KB tile (0.18 per byte) -- at 44 KB/us per SM that would need 4.1 IPC.  It is bound by its
// scalar FMA / LDS stream, not by how K arrives; the way on is QK^T and PV on mma tiles (G queries x 32 positions), not staging.
static int fa_gqa_stage_on() { static const int v = [] { const char * e = getenv("B200_FATTN_GQA_STAGE"); return e ? atoi(e) : 0; }();  return v; }
template <int KT, int VT, int D, int G>
static int fa_launch_gqa(const fa_params & p, dim3 grid, cudaStream_t st, bool want_stage) {
    cudaLaunchConfig_t lc = {};
    lc.gridDim = grid; lc.blockDim = dim3(128); lc.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = pdl_default() ? 1 : 0;
    bool staged = false;
    if constexpr (KT == T_F16) {
        if (want_stage) {                 // K and V tiles through cp.async + shared memory (f16 rows are 16-byte aligned: checked by the caller)
            constexpr size_t smem = 6 * 2 * 32 * (size_t)(2 * D + 16 + 2 * D);
            static int attr_set[16] = {0};   // per device
            int dev = 0;
            B200_CUDA_CHECK(cudaGetDevice(&dev));
            if (dev < 16 && !attr_set[dev]) {
                B200_CUDA_CHECK(cudaFuncSetAttribute(fattn_vec_gqa_kernel<KT, VT, D, G, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                attr_set[dev] = 1;
            }
            lc.dynamicSmemBytes = smem; lc.blockDim = dim3(192);
            B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, fattn_vec_gqa_kernel<KT, VT, D, G, true>, p));
            lc.dynamicSmemBytes = 0; lc.blockDim = dim3(128);
            staged = true;
        }
    }
    if (!staged) B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, fattn_vec_gqa_kernel<KT, VT, D, G>, p));
    ++g_launches;
    if (p.nsplit > 1) {
        const int64_t nrows = (int64_t)p.N * p.n_head * p.ne3;
        lc.gridDim = dim3((unsigned)((nrows * 32 + 127) / 128));
        B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, fattn_combine_kernel<D>, (const float *)p.part, p.dst, p.nbd1, nrows, p.nsplit, p.xq, p.xq_acs, p.n_head));
        ++g_launches;
    }
    return 0;
}
}  // namespace b200
#include "fattn_mma.cuh"
namespace b200 {

// warp-level tensor-core tiles for an f16 cache (fattn_mma.cuh): one CTA per SM, grid (groups of 16 queries, KV heads x ne3, splits)
template <int D, bool Q8>
static int fa_launch_mma(const fa_params & p, dim3 grid, cudaStream_t st) {
    constexpr size_t smem = fmma_smem_bytes<D, Q8>();
    static int attr_set[16] = {0};   // per device
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 16 && !attr_set[dev]) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(fattn_mma_gqa_kernel<D, Q8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set[dev] = 1;
    }
    cudaLaunchConfig_t lc = {};
    lc.gridDim = grid; lc.blockDim = dim3(32 * FMMA_NW); lc.dynamicSmemBytes = smem; lc.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = pdl_default() ? 1 : 0;
    B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, fattn_mma_gqa_kernel<D, Q8>, p));
    ++g_launches;
    if (p.nsplit > 1) {
        const int64_t nrows = (int64_t)p.N * p.n_head * p.ne3;
        lc.dynamicSmemBytes = 0; lc.blockDim = dim3(128);
        lc.gridDim = dim3((unsigned)((nrows * 32 + 127) / 128));
        B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, fattn_combine_kernel<D>, (const float *)p.part, p.dst, p.nbd1, nrows, p.nsplit, p.xq, p.xq_acs, p.n_head));
        ++g_launches;
    }
    return 0;
}
static int fa_mma_min_kv() { static const int v = [] { const char * e = getenv("B200_FATTN_MMA_MIN_KV"); return e ? atoi(e) : 512; }();  return v; }
static bool fa_mma_on() { static const bool v = [] { const char * e = getenv("B200_FATTN_MMA"); return !e || atoi(e) != 0; }();  return v; }

static int fa_gqa_min_kv() { static const int v = [] { const char * e = getenv("B200_FATTN_GQA_MIN_KV"); return e ? atoi(e) : 512; }();  return v; }

// tensor-core kernel for batches of query rows (fattn_tc.cu); returns 0 when the shape is not its to take
int fattn_tc_try(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask, const b200_tensor * dst,
                 float scale, float max_bias, float logit_softcap, void * scratch, int64_t scratch_bytes, cudaStream_t st, int * rc, void * xb = nullptr);
int64_t fattn_tc_flag_bytes(const b200_tensor * q, const b200_tensor * k);
int64_t fattn_tc_q8_scratch(const b200_tensor * q, const b200_tensor * k);
int fattn_tc_try_q8(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask, const b200_tensor * dst,
                    float scale, float max_bias, float logit_softcap, void * scratch, int64_t scratch_bytes, cudaStream_t st, int * rc, void * xb = nullptr);

}  // namespace b200

using namespace b200;

extern "C" int b200_flash_attn_supported(int k_type, int v_type, int64_t dk, int64_t dv) {
    if (dk != dv || (dk != 64 && dk != 128)) return 0;
    if (k_type != v_type) return 0;
    return k_type == T_F16 || k_type == T_Q8_0;
}

extern "C" int64_t b200_flash_attn_scratch_bytes(const b200_tensor * q, const b200_tensor * k) {
    const int64_t rows_heads = q->ne[1] * q->ne[2] * q->ne[3];
    const int ns = fa_splits(rows_heads, (int)((k->ne[1] + 31) / 32));
    int64_t part = ns > 1 ? rows_heads * ns * (2 + q->ne[0]) * 4 : 0;
    if (k->ne[1] >= fa_gqa_min_kv() && q->ne[1] <= 64) part = std::max<int64_t>(part, rows_heads * 128 * (2 + q->ne[0]) * 4);      // grouped kernel: up to 128 splits
    int64_t deq = fattn_tc_q8_scratch(q, k);                  // prefill over a q8_0 cache: f16 copies of K and V (+ mask tile flags)
    if (deq == 0) deq = fattn_tc_flag_bytes(q, k);            // tensor-core kernel: one byte per 128 x 128 mask tile
    return part > deq ? part : deq;
}

extern "C" int b200_op_flash_attn_ext(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask,
                                      const b200_tensor * dst, float scale, float max_bias, float logit_softcap,
                                      void * scratch, int64_t scratch_bytes, void * stream) {
    return b200_op_flash_attn_ext_xq(q, k, v, mask, dst, scale, max_bias, logit_softcap, scratch, scratch_bytes, nullptr, stream);
}

extern "C" int b200_flash_attn_xq_supported(const b200_tensor * q, const b200_tensor * k) {
    return q->ne[0] == 128 && q->ne[3] == 1 && q->ne[1] >= 1 && q->ne[1] <= 8 && (k->type == T_F16 || k->type == T_Q8_0);
}

// Prefill: the tensor-core kernel also stores its result as the bf16 activation rows of the output projection's MMQ (xb:
// b200_acts_bf16_bytes(D * n_head, rows), 128-byte aligned).  -1 when the tensor-core kernel does not take the shape.
extern "C" int b200_flash_attn_xb_supported(const b200_tensor * q, const b200_tensor * k, const b200_tensor * dst) {
    if (q->ne[3] != 1 || (q->ne[0] != 64 && q->ne[0] != 128) || (k->type != T_F16 && k->type != T_Q8_0)) return 0;
    if (dst->nb[1] != q->ne[0] * 4 || dst->nb[2] != dst->nb[1] * q->ne[2]) return 0;       // [D, n_head, rows] contiguous
    return fattn_tc_flag_bytes(q, k) > 0;
}
extern "C" int b200_op_flash_attn_ext_xb(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask,
                                         const b200_tensor * dst, float scale, float max_bias, float logit_softcap,
                                         void * scratch, int64_t scratch_bytes, void * xb, void * stream) {
    if (!xb || ((uintptr_t)xb & 15) || !b200_flash_attn_xb_supported(q, k, dst)) return -1;
    if (!b200_flash_attn_supported(k->type, v->type, k->ne[0], v->ne[0]) || q->type != T_F32 || dst->type != T_F32) return -1;
    if (q->ne[0] != k->ne[0] || q->nb[0] != 4 || dst->nb[0] != 4 || k->ne[1] != v->ne[1]) return -1;
    if (mask && (mask->type != T_F16 || mask->nb[0] != 2 || mask->ne[0] < k->ne[1])) return -1;
    if (q->ne[1] == 0 || k->ne[1] == 0) return -1;
    int rc = 0;
    if (fattn_tc_try(q, k, v, mask, dst, scale, max_bias, logit_softcap, scratch, scratch_bytes, (cudaStream_t)stream, &rc, xb)) return rc;
    if (k->type == T_Q8_0 && fattn_tc_try_q8(q, k, v, mask, dst, scale, max_bias, logit_softcap, scratch, scratch_bytes, (cudaStream_t)stream, &rc, xb)) return rc;
    return -1;
}

extern "C" int b200_op_flash_attn_ext_xq(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask,
                                         const b200_tensor * dst, float scale, float max_bias, float logit_softcap,
                                         void * scratch, int64_t scratch_bytes, void * xq, void * stream) {
    if (xq && !b200_flash_attn_xq_supported(q, k)) return -1;
    if (!b200_flash_attn_supported(k->type, v->type, k->ne[0], v->ne[0]) || q->type != T_F32 || dst->type != T_F32) return -1;
    if (q->ne[0] != k->ne[0] || q->nb[0] != 4 || dst->nb[0] != 4 || k->ne[1] != v->ne[1]) return -1;
    if (mask && (mask->type != T_F16 || mask->nb[0] != 2 || mask->ne[0] < k->ne[1])) return -1;
    if (k->type == T_F16 && ((((uintptr_t)k->data | k->nb[1] | k->nb[2] | k->nb[3]) & 15) || (((uintptr_t)v->data | v->nb[1] | v->nb[2] | v->nb[3]) & 7))) return -1;
    if (q->ne[1] == 0 || k->ne[1] == 0) return 0;
    {
        int rc = 0;
        if (fattn_tc_try(q, k, v, mask, dst, scale, max_bias, logit_softcap, scratch, scratch_bytes, (cudaStream_t)stream, &rc)) return rc;
        if (k->type == T_Q8_0 && fattn_tc_try_q8(q, k, v, mask, dst, scale, max_bias, logit_softcap, scratch, scratch_bytes, (cudaStream_t)stream, &rc)) return rc;
    }
    fa_params p = {};
    p.q = (const char *)q->data; p.k = (const char *)k->data; p.v = (const char *)v->data; p.mask = mask ? (const char *)mask->data : nullptr;
    p.dst = (char *)dst->data;
    p.nbq1 = q->nb[1]; p.nbq2 = q->nb[2]; p.nbq3 = q->nb[3]; p.nbk1 = k->nb[1]; p.nbk2 = k->nb[2]; p.nbk3 = k->nb[3];
    p.nbv1 = v->nb[1]; p.nbv2 = v->nb[2]; p.nbv3 = v->nb[3]; p.nbm1 = mask ? mask->nb[1] : 0; p.nbd1 = dst->nb[1];
    p.N = (int)q->ne[1]; p.n_head = (int)q->ne[2]; p.ne3 = (int)q->ne[3]; p.n_kv = (int)k->ne[1];
    p.rk2 = (int)(q->ne[2] / k->ne[2]); p.rk3 = (int)(q->ne[3] / k->ne[3]); p.rv2 = (int)(q->ne[2] / v->ne[2]); p.rv3 = (int)(q->ne[3] / v->ne[3]);
    if (logit_softcap != 0.0f) scale /= logit_softcap;
    p.scale = scale; p.max_bias = max_bias; p.softcap = logit_softcap;
    p.n_head_log2 = 1u << (unsigned)floorf(log2f((float)p.n_head));
    p.m0 = powf(2.0f, -max_bias / p.n_head_log2); p.m1 = powf(2.0f, -(max_bias / 2.0f) / p.n_head_log2);
    p.xq = (uint8_t *)xq; p.xq_acs = (uint32_t)(((q->ne[0] * q->ne[2] + 255) / 256) * ACT_ITEM);
    const int ntiles = (p.n_kv + 31) / 32;
    const int64_t rows_heads = q->ne[1] * q->ne[2] * q->ne[3];
    // long context, several queries per KV head (GQA and / or a few rows): the grouped kernel reads the cache once per group
    if (p.n_kv >= fa_gqa_min_kv() && p.N * p.rk2 >= 2 && p.rk2 == p.rv2 && p.rk3 == p.rv3 && q->ne[2] % p.rk2 == 0) {
        const int ngq = p.N * p.rk2;
        // (f16 rows must be 16-byte aligned for the copies; q8_0 rows of D = 128 / 64 are 8- / 4-byte aligned whenever the cache is)
        const uintptr_t kval = (uintptr_t)k->data | (uintptr_t)v->data | k->nb[1] | k->nb[2] | k->nb[3] | v->nb[1] | v->nb[2] | v->nb[3];
        const bool al16 = (kval & (k->type == T_F16 ? 15 : q->ne[0] == 128 ? 7 : 3)) == 0;
        if (al16 && fa_mma_on() && p.n_kv >= fa_mma_min_kv() && q->nb[1] % 16 == 0 && q->nb[2] % 16 == 0 && q->nb[3] % 16 == 0 &&
            ((uintptr_t)q->data & 15) == 0 && (!mask || ((((uintptr_t)mask->data | mask->nb[1]) & 15) == 0))) {
            // f16 cache, long context: m16n8k16 tiles, the whole group (up to 16 queries) per CTA, one CTA per SM, one wave
            const int nx = (ngq + 15) / 16;
            const int64_t ctas = (int64_t)nx * (q->ne[2] / p.rk2) * q->ne[3];
            int ns = (int)std::min<int64_t>(std::max<int64_t>(1, 148 / ctas), std::max(1, ntiles / 4));
            ns = std::max(1, std::min(ns, 128));
            if (ns > 1 && (!scratch || scratch_bytes < rows_heads * ns * (2 + q->ne[0]) * 4)) ns = 1;
            if (!(xq && ns == 1)) {
                p.nsplit = ns;
                p.tiles_per_split = (ntiles + ns - 1) / ns;
                p.part = (float *)scratch;
                const dim3 grid((unsigned)nx, (unsigned)((q->ne[2] / p.rk2) * q->ne[3]), (unsigned)ns);
                cudaStream_t st = (cudaStream_t)stream;
                if (k->type == T_F16) return q->ne[0] == 128 ? fa_launch_mma<128, false>(p, grid, st) : fa_launch_mma<64, false>(p, grid, st);
                return q->ne[0] == 128 ? fa_launch_mma<128, true>(p, grid, st) : fa_launch_mma<64, true>(p, grid, st);
            }
        }
        const int G = ngq <= 4 ? 4 : 8;
        const int nx = (ngq + G - 1) / G;
        const int64_t ctas = (int64_t)nx * (q->ne[2] / p.rk2) * q->ne[3];
        // f16 cache at long context: the staged form, CTAs of 6 warps that own an SM each, one wave of them
        static const int stage_min_kv = [] { const char * e = getenv("B200_FATTN_GQA_STAGE_MIN_KV"); return e ? atoi(e) : 4096; }();
        const bool staged = k->type == T_F16 && fa_gqa_stage_on() && p.n_kv >= stage_min_kv;
        int ns = staged ? (int)std::min<int64_t>(std::max<int64_t>(1, (148 + ctas - 1) / ctas), std::max(1, ntiles / 12))
                        : (int)std::min<int64_t>(std::max<int64_t>(1, (4 * 148 + ctas - 1) / ctas), std::max(1, ntiles / 4));
        ns = std::max(1, std::min(ns, 128));
        if (ns > 1 && (!scratch || scratch_bytes < rows_heads * ns * (2 + q->ne[0]) * 4)) ns = 1;
        if (xq && ns == 1) return -1;              // (the grouped kernel writes activations through the combine kernel only)
        p.nsplit = ns;
        p.tiles_per_split = (ntiles + ns - 1) / ns;
        p.part = (float *)scratch;
        const dim3 grid((unsigned)nx, (unsigned)((q->ne[2] / p.rk2) * q->ne[3]), (unsigned)ns);
        cudaStream_t st = (cudaStream_t)stream;
        const int D = (int)q->ne[0];
        const bool f16 = k->type == T_F16;
        if (G == 4) {
            if (D == 128) return f16 ? fa_launch_gqa<T_F16, T_F16, 128, 4>(p, grid, st, staged) : fa_launch_gqa<T_Q8_0, T_Q8_0, 128, 4>(p, grid, st, false);
            return f16 ? fa_launch_gqa<T_F16, T_F16, 64, 4>(p, grid, st, staged) : fa_launch_gqa<T_Q8_0, T_Q8_0, 64, 4>(p, grid, st, false);
        }
        if (D == 128) return f16 ? fa_launch_gqa<T_F16, T_F16, 128, 8>(p, grid, st, staged) : fa_launch_gqa<T_Q8_0, T_Q8_0, 128, 8>(p, grid, st, false);
        return f16 ? fa_launch_gqa<T_F16, T_F16, 64, 8>(p, grid, st, staged) : fa_launch_gqa<T_Q8_0, T_Q8_0, 64, 8>(p, grid, st, false);
    }
    p.nsplit = fa_splits(rows_heads, ntiles);
    if (p.nsplit > 1 && (!scratch || scratch_bytes < rows_heads * p.nsplit * (2 + q->ne[0]) * 4)) p.nsplit = 1;
    p.tiles_per_split = (ntiles + p.nsplit - 1) / p.nsplit;
    p.part = (float *)scratch;
    const dim3 grid((unsigned)q->ne[1], (unsigned)(q->ne[2] * q->ne[3]), (unsigned)p.nsplit);
    cudaStream_t st = (cudaStream_t)stream;
    const int D = (int)q->ne[0];
    if (k->type == T_F16) {
        // f16 cache with 16-byte aligned rows: K / V tiles staged through shared memory by bulk copies
        const bool al16 = ((((uintptr_t)v->data | v->nb[1] | v->nb[2] | v->nb[3]) & 15) == 0);
        if (fa_stage_on() && al16) return D == 128 ? fa_launch_stage<128>(p, grid, st) : fa_launch_stage<64>(p, grid, st);
        return D == 128 ? fa_launch<T_F16, T_F16, 128>(p, grid, st) : fa_launch<T_F16, T_F16, 64>(p, grid, st);
    }
    return D == 128 ? fa_launch<T_Q8_0, T_Q8_0, 128>(p, grid, st) : fa_launch<T_Q8_0, T_Q8_0, 64>(p, grid, st);
}
