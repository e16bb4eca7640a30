// mmvq_units.cuh -- per-format "unit" arithmetic of the decode GEMV.
//
// A UNIT is 32 weights of one row that one thread multiplies with 32 activations it keeps in
// registers.  Every format has 8 units per 256 weights, so a row of K weights has K/32 units:
//   Q4_0/Q8_0 : unit u = block u                      -> 1 group of 32 consecutive elements
//   Q4_K/Q5_K : unit u = 16-byte chunk c=u%8 of qs    -> 2 groups of 16 (low / high nibbles)
//   Q6_K      : unit u = (half h, 8 lanes l0) of ql/qh -> 4 groups of 8
// Activations are quantised to q8_1 semantics (d = amax/127, q = roundf(x/d); reference
// ggml-cuda/quantize.cu:4-49) directly into registers -- the fused prologue of north-star item (2).
// The integer dot products are those of reference ggml-cuda/vecdotq.cuh:25-44 (q4_0), 153-165
// (q8_0), 359-381 (q4_K), 415-444 (q5_K), 478-498 (q6_K); they are order independent and therefore
// bit-exact; the fp32 combination order is ours.
#pragma once
#include "common.cuh"

namespace b200 {

template <int TYPE> struct unit_traits;
template <> struct unit_traits<T_Q4_0> { static constexpr int G = 1; };
template <> struct unit_traits<T_Q8_0> { static constexpr int G = 1; };
template <> struct unit_traits<T_Q4_K> { static constexpr int G = 2; };
template <> struct unit_traits<T_Q5_K> { static constexpr int G = 2; };
template <> struct unit_traits<T_Q6_K> { static constexpr int G = 4; };

// element index (within the row) of the first element of group g of unit u
template <int TYPE> __device__ __forceinline__ int unit_group_elem0(int u, int g) {
    if constexpr (TYPE == T_Q4_0 || TYPE == T_Q8_0) { return 32 * u; }
    else if constexpr (TYPE == T_Q4_K || TYPE == T_Q5_K) {
        const int b = u >> 3, c = u & 7;
        return 256 * b + 64 * (c >> 1) + 16 * (c & 1) + 32 * g;
    } else {  // Q6_K
        const int b = u >> 3, c = u & 7;
        return 256 * b + 128 * (c >> 2) + 8 * (c & 3) + 32 * g;
    }
}

// quantised activations of one unit for one activation column
template <int G> struct act_regs {
    int   q[8];     // 32 int8, group-major
    float d8[G];    // q8_1 scale of the 32-block each group belongs to
    int   sq[G];    // sum of this unit's q inside each group (exact int; the "dot2" of vecdotq.cuh:376)
};

// Quantise the 32 activations of unit u (column pointer x, K valid elements) into registers.
// Threads of the same q8_1 block are G adjacent lanes, so amax is completed with xor-shuffles.
// Must be called by all 32 lanes (inactive units pass valid=false and contribute zeros).
template <int TYPE> __device__ __forceinline__ void quantize_unit(
        const float * __restrict__ x, int u, bool valid, bool vec_ok, act_regs<unit_traits<TYPE>::G> & a) {
    constexpr int G = unit_traits<TYPE>::G, E = 32 / G;   // elements per group
#pragma unroll
    for (int g = 0; g < G; ++g) {
        float v[E];
        if (valid) {
            const float * src = x + unit_group_elem0<TYPE>(u, g);
            if (vec_ok) {
#pragma unroll
                for (int i = 0; i < E / 4; ++i) {
                    const float4 t = __ldg(reinterpret_cast<const float4 *>(src) + i);
                    v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
                }
            } else {
#pragma unroll
                for (int i = 0; i < E; ++i) v[i] = __ldg(src + i);
            }
        } else {
#pragma unroll
            for (int i = 0; i < E; ++i) v[i] = 0.0f;
        }
        float amax = 0.0f;
#pragma unroll
        for (int i = 0; i < E; ++i) amax = fmaxf(amax, fabsf(v[i]));
#pragma unroll
        for (int o = 1; o < G; o <<= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        const float d = __fdiv_rn(amax, 127.0f);
        int sq = 0;
#pragma unroll
        for (int i = 0; i < E / 4; ++i) {
            int packed = 0;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int qi = amax == 0.0f ? 0 : (int)roundf(__fdiv_rn(v[4 * i + j], d));
                sq += qi;
                packed |= (qi & 0xFF) << (8 * j);
            }
            a.q[g * (E / 4) + i] = packed;
        }
        a.d8[g] = __half2float(__float2half_rn(d));   // q8_1 stores d as binary16 (reference quantize.cu:47)
        a.sq[g] = sq;
    }
}

// exact int -> float for |i| < 2^22 on the FMA/ALU pipes (avoids the quarter-rate I2F path)
__device__ __forceinline__ float i2f_small(int i) {
    return __int_as_float(0x4B400000 + i) - 12582912.0f;
}

// 6-bit (scale,min) pairs of sub-blocks 2j and 2j+1 of a K-quant header (reference
// ggml-quants.c:625-632), both at once: returns sc = {sc[2j], sc[2j+1]} and mn likewise as bytes
// 0 and 1 of the outputs.  w0,w1,w2 are the three 32-bit words of scales[12].
__device__ __forceinline__ void k4_scales_pair(uint32_t w0, uint32_t w1, uint32_t w2, int j, uint32_t & sc2, uint32_t & mn2) {
    const int sh = 16 * (j & 1);
    const uint32_t h0 = (w0 >> sh) & 0xFFFFu, h1 = (w1 >> sh) & 0xFFFFu, h2 = (w2 >> sh) & 0xFFFFu;
    const uint32_t sc_lo = h0 & 0x3F3Fu, mn_lo = h1 & 0x3F3Fu;
    const uint32_t sc_hi = (h2 & 0x0F0Fu) | ((h0 >> 2) & 0x3030u);
    const uint32_t mn_hi = ((h2 >> 4) & 0x0F0Fu) | ((h1 >> 2) & 0x3030u);
    sc2 = j < 2 ? sc_lo : sc_hi;
    mn2 = j < 2 ? mn_lo : mn_hi;
}

// Result of the integer part of one unit x one column, kept so that the parity tests can read
// the exact int32 sums out of the very code the production kernel runs.
template <int G> struct unit_ints { int dot[G]; };

// ---- the per-format unit dot products -------------------------------------------------------
// rowp: shared-memory byte address of the row (2-byte aligned; 16-byte aligned when A16).
// Adds this unit's contribution for M columns to acc[M]; optionally exports the int sums.
template <int TYPE, int M, bool A16>
__device__ __forceinline__ void unit_dot(uint32_t rowp, int u, const act_regs<unit_traits<TYPE>::G> (&a)[M],
                                         float (&acc)[M], unit_ints<unit_traits<TYPE>::G> * ints = nullptr) {
    if constexpr (TYPE == T_Q4_0) {
        const uint32_t bp = rowp + 18u * u;
        int w[4];
        const float d = half_bits_to_float(lds16(bp));
        lds_words_a2<4>(bp + 2, w);
        int lo[4], hi[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { lo[i] = w[i] & 0x0F0F0F0F; hi[i] = (w[i] >> 4) & 0x0F0F0F0F; }
#pragma unroll
        for (int m = 0; m < M; ++m) {
            int s = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) { s = __dp4a(lo[i], a[m].q[i], s); s = __dp4a(hi[i], a[m].q[4 + i], s); }
            if (ints) ints[m].dot[0] = s;
            acc[m] = fmaf(d * a[m].d8[0], i2f_small(s - 8 * a[m].sq[0]), acc[m]);
        }
    } else if constexpr (TYPE == T_Q8_0) {
        const uint32_t bp = rowp + 34u * u;
        int w[8];
        const float d = half_bits_to_float(lds16(bp));
        lds_words_a2<8>(bp + 2, w);
#pragma unroll
        for (int m = 0; m < M; ++m) {
            int s = 0;
#pragma unroll
            for (int i = 0; i < 8; ++i) s = __dp4a(w[i], a[m].q[i], s);
            if (ints) ints[m].dot[0] = s;
            acc[m] = fmaf(d * a[m].d8[0], i2f_small(s), acc[m]);
        }
    } else if constexpr (TYPE == T_Q4_K || TYPE == T_Q5_K) {
        constexpr int BB = fmt<TYPE>::BB, QS0 = (TYPE == T_Q4_K) ? 16 : 48;
        const int c = u & 7, j = c >> 1;
        const uint32_t bp = rowp + (uint32_t)BB * (u >> 3);
        int hd[4], w[4];
        if constexpr (A16) {
            const int4 t = lds128(bp); hd[0] = t.x; hd[1] = t.y; hd[2] = t.z; hd[3] = t.w;
            const int4 q = lds128(bp + QS0 + 16 * c); w[0] = q.x; w[1] = q.y; w[2] = q.z; w[3] = q.w;
        } else {
            lds_words_a2<4>(bp, hd);
            lds_words_a2<4>(bp + QS0 + 16 * c, w);
        }
        int lo[4], hi[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { lo[i] = w[i] & 0x0F0F0F0F; hi[i] = (w[i] >> 4) & 0x0F0F0F0F; }
        if constexpr (TYPE == T_Q5_K) {
            int qh[4];
            if constexpr (A16) { const int4 t = lds128(bp + 16 + 16 * (c & 1)); qh[0] = t.x; qh[1] = t.y; qh[2] = t.z; qh[3] = t.w; }
            else { lds_words_a2<4>(bp + 16 + 16 * (c & 1), qh); }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                lo[i] |= ((qh[i] >> (2 * j)) & 0x01010101) << 4;
                hi[i] |= ((qh[i] >> (2 * j + 1)) & 0x01010101) << 4;
            }
        }
        uint32_t sc2, mn2;
        k4_scales_pair((uint32_t)hd[1], (uint32_t)hd[2], (uint32_t)hd[3], j, sc2, mn2);
        const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&hd[0]));
        const int sc0 = sc2 & 0xFF, sc1 = sc2 >> 8, mn0 = mn2 & 0xFF, mn1 = mn2 >> 8;
#pragma unroll
        for (int m = 0; m < M; ++m) {
            int s0 = 0, s1 = 0;
#pragma unroll
            for (int i = 0; i < 4; ++i) { s0 = __dp4a(lo[i], a[m].q[i], s0); s1 = __dp4a(hi[i], a[m].q[4 + i], s1); }
            if (ints) { ints[m].dot[0] = s0; ints[m].dot[1] = s1; }
            const float pd = a[m].d8[0] * i2f_small(s0 * sc0) + a[m].d8[1] * i2f_small(s1 * sc1);
            const float pm = a[m].d8[0] * i2f_small(a[m].sq[0] * mn0) + a[m].d8[1] * i2f_small(a[m].sq[1] * mn1);
            acc[m] += dm.x * pd - dm.y * pm;
        }
    } else {  // Q6_K
        const int c = u & 7, h = c >> 2, l0 = 8 * (c & 3);
        const uint32_t bp = rowp + 210u * (u >> 3);
        int qa[2], qb[2], qh[2];
        lds_words_a2<2>(bp + 64 * h + l0, qa);
        lds_words_a2<2>(bp + 64 * h + 32 + l0, qb);
        lds_words_a2<2>(bp + 128 + 32 * h + l0, qh);
        // scales[8h + (l0>>4) + {0,2,4,6}] : 7 consecutive bytes starting at an arbitrary byte
        const uint32_t sa = bp + 192 + 8 * h + (l0 >> 4);
        const uint32_t sbase = sa & ~3u, ssh = (sa & 3u) * 8u;
        const uint32_t s0w = lds32(sbase), s1w = lds32(sbase + 4), s2w = lds32(sbase + 8);
        const uint32_t sw0 = __funnelshift_r(s0w, s1w, ssh), sw1 = __funnelshift_r(s1w, s2w, ssh);
        const int sc[4] = { (int)(int8_t)(sw0 & 0xFF), (int)(int8_t)((sw0 >> 16) & 0xFF),
                            (int)(int8_t)(sw1 & 0xFF), (int)(int8_t)((sw1 >> 16) & 0xFF) };
        const float d = half_bits_to_float(lds16(bp + 208));
        int v[4][2];
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            v[0][i] = (qa[i] & 0x0F0F0F0F)        | ((qh[i] << 4) & 0x30303030);
            v[1][i] = (qb[i] & 0x0F0F0F0F)        | ((qh[i] << 2) & 0x30303030);
            v[2][i] = ((qa[i] >> 4) & 0x0F0F0F0F) | ( qh[i]       & 0x30303030);
            v[3][i] = ((qb[i] >> 4) & 0x0F0F0F0F) | ((qh[i] >> 2) & 0x30303030);
        }
#pragma unroll
        for (int m = 0; m < M; ++m) {
            float p = 0.0f;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                // sum (v-32)*q = sum v*q - 32*sum q : same int32 as the __vsubss4 form (vecdotq.cuh:492)
                int s = __dp4a(v[g][0], a[m].q[2 * g], __dp4a(v[g][1], a[m].q[2 * g + 1], -32 * a[m].sq[g]));
                if (ints) ints[m].dot[g] = s;
                p = fmaf(a[m].d8[g], i2f_small(s * sc[g]), p);
            }
            acc[m] = fmaf(d, p, acc[m]);
        }
    }
}

}  // namespace b200
