// mmvq_units.cuh -- per-format arithmetic of the decode GEMV (v2: one 256-weight ITEM per thread).
//
// An ITEM is 256 consecutive weights of one row: one K-quant super-block (Q4_K 144 B, Q5_K 176 B,
// Q6_K 210 B) or eight consecutive 32-blocks (Q4_0 144 B, Q8_0 272 B; the last item of a row may
// hold fewer when K % 256 != 0).  One thread multiplies one item with the matching 256 activations
// of M columns, so the K-quant header (fp16 scales, 6-bit sub-scales) is unpacked once per 256
// weights, and 144/176/272-byte items are read with conflict-free LDS.128 (odd multiples of 16 B).
//
// Activations are quantised to q8_1 semantics (d = amax/127, q = roundf(x/d); reference
// ggml-cuda/quantize.cu:4-49) ONCE per CTA into shared memory -- the fused prologue of north-star
// item (2) -- in this layout, per column, per 32-block: 48 bytes
//     q[32] int8 | d8 (fp32 value of the fp16-rounded scale) | dsq | m32a | m32b
// and 400 bytes per item (8 x 48 + 16 pad: an odd multiple of 16 B, so that lanes reading
// consecutive items with LDS.128 hit distinct banks).  The tail serves every weight format:
//     dsq  = d8 * sum q (fp32)            Q4_K, Q5_K: the dmin term (vecdotq.cuh:376,437)
//     m32a = -32 * sum q[0..15]  (int)    Q6_K: folds __vsubss4(., 32) (vecdotq.cuh:492)
//     m32b = -32 * sum q[16..31] (int)    Q4_0: (m32a + m32b) >> 2 = -8 * sum q (vecdotq.cuh:43)
// The same layout is produced in global memory by quantize_acts_kernel (quantize.cu) for GEMVs that
// take pre-quantised activations (large K, small N: per-CTA re-quantisation would cost more than
// the dot products there).
// The integer dot products are those of reference ggml-cuda/vecdotq.cuh:25-44 (q4_0), 153-165
// (q8_0), 359-381 (q4_K), 415-444 (q5_K), 478-498 (q6_K), 49-80 (q4_1), 82-115 (q5_0), 117-150 (q5_1); integer sums are order
// independent and therefore bit-exact; the fp32 combination order is ours.  Where the reference multiplies a block minimum or
// offset with s8 = fp16(sum of the unquantised x) (q4_0, q4_1, q5_0, q5_1) we use the exact d8 * sum q of the layout's tail.
#pragma once
#include "common.cuh"

namespace b200 {

template <int TYPE> struct item_traits;
template <> struct item_traits<T_Q4_0> { static constexpr int BYTES = 144; };
template <> struct item_traits<T_Q8_0> { static constexpr int BYTES = 272; };
template <> struct item_traits<T_Q2_K> { static constexpr int BYTES = 84; };    // scales[16] | qs[64] | d | dmin
template <> struct item_traits<T_Q3_K> { static constexpr int BYTES = 110; };   // hmask[32] | qs[64] | scales[12] | d
template <> struct item_traits<T_IQ4_NL> { static constexpr int BYTES = 144; };  // 8 x {d, qs[16]}: 4-bit indices into a fixed table
template <> struct item_traits<T_IQ4_XS> { static constexpr int BYTES = 136; };  // d | scales_h | scales_l[4] | qs[128]
template <> struct item_traits<T_Q4_1> { static constexpr int BYTES = 160; };   // 8 x {d, m, qs[16]}
template <> struct item_traits<T_Q5_0> { static constexpr int BYTES = 176; };   // 8 x {d, qh[4], qs[16]}
template <> struct item_traits<T_Q5_1> { static constexpr int BYTES = 192; };   // 8 x {d, m, qh[4], qs[16]}
template <> struct item_traits<T_Q4_K> { static constexpr int BYTES = 144; };
template <> struct item_traits<T_Q5_K> { static constexpr int BYTES = 176; };
template <> struct item_traits<T_Q6_K> { static constexpr int BYTES = 210; };

constexpr int ACT_BLK = 48;      // bytes per 32-block of quantised activations in shared memory
constexpr int ACT_ITEM = 400;    // bytes per 256-element item (8 * 48 + 16)

// exact int -> float for |i| < 2^22 on the FMA/ALU pipes (avoids the quarter-rate I2F path)
__device__ __forceinline__ float i2f_small(int i) {
    return __int_as_float(0x4B400000 + i) - 12582912.0f;
}
// byte k of x -> float, exact (PRMT + FADD)
__device__ __forceinline__ float byte2f(uint32_t x, int k) {
    return __uint_as_float(__byte_perm(x, 0x4B000000u, 0x7440 + k)) - 8388608.0f;
}
// unsigned bytes x signed bytes dot product (the x16 nibble trick needs u8 weights)
__device__ __forceinline__ int dp4a_us(uint32_t a, int b, int c) {
    int d; asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}
__device__ __forceinline__ int dp4a_ss(int a, int b, int c) { return __dp4a(a, b, c); }

// N consecutive 32-bit words from shared memory; A16: address is 16-byte aligned and N % 4 == 0
template <int N, bool A16> __device__ __forceinline__ void lds_words(uint32_t saddr, int (&w)[N]) {
    if constexpr (A16 && N % 4 == 0) {
#pragma unroll
        for (int i = 0; i < N / 4; ++i) {
            const int4 t = lds128(saddr + 16 * i);
            w[4 * i] = t.x; w[4 * i + 1] = t.y; w[4 * i + 2] = t.z; w[4 * i + 3] = t.w;
        }
    } else {
        lds_words_a2<N>(saddr, w);
    }
}

// activations of one 32-block of one column: 8 words of int8 + (d8, aux0, aux1)
struct act_blk { int q[8]; float d8, dsq; int m32a, m32b; };
__device__ __forceinline__ void lds_act(uint32_t ap, act_blk & a) {
    const int4 t0 = lds128(ap), t1 = lds128(ap + 16), t2 = lds128(ap + 32);
    a.q[0] = t0.x; a.q[1] = t0.y; a.q[2] = t0.z; a.q[3] = t0.w;
    a.q[4] = t1.x; a.q[5] = t1.y; a.q[6] = t1.z; a.q[7] = t1.w;
    a.d8 = __int_as_float(t2.x); a.dsq = __int_as_float(t2.y); a.m32a = t2.z; a.m32b = t2.w;
}

// halfword h (static index) of a register array of words
template <int NW> __device__ __forceinline__ uint32_t halfword(const int (&w)[NW], int h) {
    return (h & 1) ? ((uint32_t)w[h >> 1] >> 16) : ((uint32_t)w[h >> 1] & 0xFFFFu);
}
// 32-bit word starting at halfword h (static index)
template <int NW> __device__ __forceinline__ int word_at_half(const int (&w)[NW], int h) {
    return (h & 1) ? (int)__funnelshift_r((uint32_t)w[h >> 1], (uint32_t)w[(h >> 1) + 1], 16) : w[h >> 1];
}

// ---- the per-format item dot products ---------------------------------------------------------
// wp: shared byte address of the item (2-byte aligned; 16-byte aligned when A16)
// ap: shared byte address of the item's activations of column 0; acs: bytes between columns
// nblk: valid 32-blocks in this item (8 except the tail item of Q4_0/Q8_0 rows)
// dbg: optional global pointer to this item's 8 int32 block sums of column 0; dms: ints between columns
template <int TYPE, int M, bool A16, bool DBG>
__device__ __forceinline__ void item_dot(uint32_t wp, uint32_t ap, uint32_t acs, int nblk, float (&acc)[M],
                                         int32_t * dbg, int64_t dms) {
    if constexpr (TYPE == T_Q4_0) {
        int w[36];
        lds_words<36, A16>(wp, w);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (i < nblk) {
                const float dw = half_bits_to_float(halfword<36>(w, 9 * i));
                uint32_t lo[4], hi[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint32_t q = (uint32_t)word_at_half<36>(w, 9 * i + 1 + 2 * j);
                    lo[j] = q & 0x0F0F0F0Fu; hi[j] = q & 0xF0F0F0F0u;
                }
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    act_blk a; lds_act(ap + m * acs + i * ACT_BLK, a);
                    int s0 = 0, s1 = 0;
#pragma unroll
                    for (int j = 0; j < 4; ++j) { s0 = dp4a_us(lo[j], a.q[j], s0); s1 = dp4a_us(hi[j], a.q[4 + j], s1); }
                    const int s = s0 + (s1 >> 4);
                    if constexpr (DBG) dbg[m * dms + i] = s;
                    acc[m] = fmaf(dw * a.d8, i2f_small(s + ((a.m32a + a.m32b) >> 2)), acc[m]);
                }
            }
        }
    } else if constexpr (TYPE == T_Q4_1) {
        int w[40];                                       // block i = words 5i .. 5i+4: (d, m) | qs[16]
        lds_words<40, A16>(wp, w);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (i < nblk) {
                const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&w[5 * i]));
                uint32_t lo[4], hi[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) { lo[j] = (uint32_t)w[5 * i + 1 + j] & 0x0F0F0F0Fu; hi[j] = (uint32_t)w[5 * i + 1 + j] & 0xF0F0F0F0u; }
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    act_blk a; lds_act(ap + m * acs + i * ACT_BLK, a);
                    int s0 = 0, s1 = 0;
#pragma unroll
                    for (int j = 0; j < 4; ++j) { s0 = dp4a_us(lo[j], a.q[j], s0); s1 = dp4a_us(hi[j], a.q[4 + j], s1); }
                    const int s = s0 + (s1 >> 4);
                    if constexpr (DBG) dbg[m * dms + i] = s;
                    acc[m] = fmaf(dm.x * a.d8, i2f_small(s), fmaf(dm.y, a.dsq, acc[m]));
                }
            }
        }
    } else if constexpr (TYPE == T_Q5_0 || TYPE == T_Q5_1) {
        // fifth bits: bit j of qh belongs to element j; the four bits of a word of low (high) nibbles are spread to bit 4 of its
        // bytes by one multiply (x * 0x00204081 puts bit k of x at bit 8k: the four shifted copies do not overlap)
        constexpr int NW = TYPE == T_Q5_0 ? 44 : 48;
        int w[NW];
        lds_words<NW, A16>(wp, w);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (i < nblk) {
                float dw, mw = 0.0f; uint32_t qh, qs[4];
                if constexpr (TYPE == T_Q5_0) {          // block i = halfwords 11i .. 11i+10: d | qh[4] | qs[16]
                    dw = half_bits_to_float(halfword<NW>(w, 11 * i));
                    qh = (uint32_t)word_at_half<NW>(w, 11 * i + 1);
#pragma unroll
                    for (int j = 0; j < 4; ++j) qs[j] = (uint32_t)word_at_half<NW>(w, 11 * i + 3 + 2 * j);
                } else {                                 // block i = words 6i .. 6i+5: (d, m) | qh | qs[16]
                    const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&w[6 * i]));
                    dw = dm.x; mw = dm.y;
                    qh = (uint32_t)w[6 * i + 1];
#pragma unroll
                    for (int j = 0; j < 4; ++j) qs[j] = (uint32_t)w[6 * i + 2 + j];
                }
                uint32_t lo[4], hi[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const uint32_t b0 = ((((qh >> (4 * j)) & 0xFu) * 0x00204081u) & 0x01010101u) << 4;
                    const uint32_t b1 = ((((qh >> (16 + 4 * j)) & 0xFu) * 0x00204081u) & 0x01010101u) << 4;
                    lo[j] = (qs[j] & 0x0F0F0F0Fu) | b0;
                    hi[j] = ((qs[j] >> 4) & 0x0F0F0F0Fu) | b1;
                }
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    act_blk a; lds_act(ap + m * acs + i * ACT_BLK, a);
                    int s = 0;
#pragma unroll
                    for (int j = 0; j < 4; ++j) { s = dp4a_us(lo[j], a.q[j], s); s = dp4a_us(hi[j], a.q[4 + j], s); }
                    if constexpr (DBG) dbg[m * dms + i] = s;
                    if constexpr (TYPE == T_Q5_0) acc[m] = fmaf(dw * a.d8, i2f_small(s + ((a.m32a + a.m32b) >> 1)), acc[m]);   // (q - 16): -16 * sum q
                    else acc[m] = fmaf(dw * a.d8, i2f_small(s), fmaf(mw, a.dsq, acc[m]));
                }
            }
        }
    } else if constexpr (TYPE == T_Q8_0) {
        int w[68];
        lds_words<68, A16>(wp, w);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (i < nblk) {
                const float dw = half_bits_to_float(halfword<68>(w, 17 * i));
                int q[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) q[j] = word_at_half<68>(w, 17 * i + 1 + 2 * j);
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    act_blk a; lds_act(ap + m * acs + i * ACT_BLK, a);
                    int s = 0;
#pragma unroll
                    for (int j = 0; j < 8; ++j) s = dp4a_ss(q[j], a.q[j], s);
                    if constexpr (DBG) dbg[m * dms + i] = s;
                    acc[m] = fmaf(dw * a.d8, i2f_small(s), acc[m]);
                }
            }
        }
    } else if constexpr (TYPE == T_IQ4_NL) {
        int w[36];                                       // block i = halfwords 9i .. 9i+8: d | qs[16]
        lds_words<36, A16>(wp, w);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            if (i < nblk) {
                const float dw = half_bits_to_float(halfword<36>(w, 9 * i));
                int vlo[4], vhi[4];
#pragma unroll
                for (int j = 0; j < 4; ++j) iq4_lookup8((uint32_t)word_at_half<36>(w, 9 * i + 1 + 2 * j), vlo[j], vhi[j]);
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    act_blk a; lds_act(ap + m * acs + i * ACT_BLK, a);
                    int s = 0;
#pragma unroll
                    for (int j = 0; j < 4; ++j) { s = dp4a_ss(vlo[j], a.q[j], s); s = dp4a_ss(vhi[j], a.q[4 + j], s); }
                    if constexpr (DBG) dbg[m * dms + i] = s;
                    acc[m] = fmaf(dw * a.d8, i2f_small(s), acc[m]);
                }
            }
        }
    } else if constexpr (TYPE == T_IQ4_XS) {
        int hd[2];
        lds_words_a2<2>(wp, hd);                         // d | scales_h, scales_l[4]
        const float d = half_bits_to_float((uint32_t)hd[0] & 0xFFFFu);
        const uint32_t sh = (uint32_t)hd[0] >> 16, sl = (uint32_t)hd[1];
        float accx[M];
#pragma unroll
        for (int m = 0; m < M; ++m) accx[m] = 0.0f;
#pragma unroll
        for (int ib = 0; ib < 8; ++ib) {
            int q[4], vlo[4], vhi[4];
            lds_words_a2<4>(wp + 8 + 16 * ib, q);
            const int ls = (int)(((sl >> (4 * ib)) & 0xFu) | (((sh >> (2 * ib)) & 3u) << 4)) - 32;      // scales_l[ib / 2] >> 4 (ib % 2)
#pragma unroll
            for (int j = 0; j < 4; ++j) iq4_lookup8((uint32_t)q[j], vlo[j], vhi[j]);
#pragma unroll
            for (int m = 0; m < M; ++m) {
                act_blk a; lds_act(ap + m * acs + ib * ACT_BLK, a);
                int s = 0;
#pragma unroll
                for (int j = 0; j < 4; ++j) { s = dp4a_ss(vlo[j], a.q[j], s); s = dp4a_ss(vhi[j], a.q[4 + j], s); }
                if constexpr (DBG) dbg[m * dms + ib] = s;
                accx[m] = fmaf(a.d8 * (float)ls, i2f_small(s), accx[m]);
            }
        }
#pragma unroll
        for (int m = 0; m < M; ++m) acc[m] = fmaf(d, accx[m], acc[m]);
    } else if constexpr (TYPE == T_Q2_K) {
        // vecdotq.cuh:218-243.  32-block 4n + j: bits 2j..2j+1 of qs[32n .. 32n+31], a 4-bit scale / 4-bit min pair per 16 elements
        int sc[4], q[8], dmw[1];
        lds_words_a2<4>(wp, sc);
        lds_words_a2<1>(wp + 80, dmw);
        const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&dmw[0]));
        float accd[M], accm[M];
#pragma unroll
        for (int m = 0; m < M; ++m) { accd[m] = 0.0f; accm[m] = 0.0f; }
#pragma unroll
        for (int n = 0; n < 2; ++n) {
            lds_words_a2<8>(wp + 16 + 32 * n, q);           // qs[32n .. 32n+31]: every byte serves the four j
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const uint32_t scw = (uint32_t)sc[2 * n + (j >> 1)] >> (16 * (j & 1));       // bytes 8n + 2j, 8n + 2j + 1
                const int sa = scw & 0xF, ma = (scw >> 4) & 0xF, sb = (scw >> 8) & 0xF, mb = (scw >> 12) & 0xF;
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    act_blk a; lds_act(ap + m * acs + (4 * n + j) * ACT_BLK, a);
                    int s0 = 0, s1 = 0;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        s0 = dp4a_us(((uint32_t)q[k] >> (2 * j)) & 0x03030303u, a.q[k], s0);
                        s1 = dp4a_us(((uint32_t)q[4 + k] >> (2 * j)) & 0x03030303u, a.q[4 + k], s1);
                    }
                    if constexpr (DBG) dbg[m * dms + 4 * n + j] = s0 + s1;
                    accd[m] = fmaf(a.d8, i2f_small(s0 * sa + s1 * sb), accd[m]);
                    accm[m] = fmaf(a.d8, i2f_small(-(a.m32a >> 5) * ma - (a.m32b >> 5) * mb), accm[m]);      // m32a = -32 * sum of the first 16 q8
                }
            }
        }
#pragma unroll
        for (int m = 0; m < M; ++m) acc[m] += dm.x * accd[m] - dm.y * accm[m];
    } else if constexpr (TYPE == T_Q3_K) {
        // vecdotq.cuh:301-331.  code = 2 low bits | hmask bit << 2, minus 4 (folded: -4 * sum q8 = m32 >> 3); 6-bit scales minus 32
        int hm[8], q[8], scw[3];
        lds_words_a2<8>(wp, hm);
        lds_words_a2<3>(wp + 96, scw);
        const float d = half_bits_to_float(lds16(wp + 108));
        float acc3[M];
#pragma unroll
        for (int m = 0; m < M; ++m) acc3[m] = 0.0f;
#pragma unroll
        for (int n = 0; n < 2; ++n) {
            lds_words_a2<8>(wp + 32 + 32 * n, q);
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                // scale indices 8n + 2j, 8n + 2j + 1: low nibble from scales[is % 8] (>> 4 for is >= 8), high pair from scales[8 + is % 4] >> 2 (is / 4)
                int scv[2];
#pragma unroll
                for (int h = 0; h < 2; ++h) {
                    const int is = 8 * n + 2 * j + h, b = is & 3, g = is >> 2;
                    const uint32_t lo = ((uint32_t)scw[(g & 1)] >> (8 * b + (g >= 2 ? 4 : 0))) & 0xFu;
                    const uint32_t hi = ((uint32_t)scw[2] >> (8 * b + 2 * g)) & 3u;
                    scv[h] = (int)(lo | (hi << 4)) - 32;
                }
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    act_blk a; lds_act(ap + m * acs + (4 * n + j) * ACT_BLK, a);
                    int s0 = a.m32a >> 3, s1 = a.m32b >> 3;      // -4 * sum q8 of each half
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const uint32_t v0 = (((uint32_t)q[k] >> (2 * j)) & 0x03030303u) | ((((uint32_t)hm[k] >> (4 * n + j)) & 0x01010101u) << 2);
                        const uint32_t v1 = (((uint32_t)q[4 + k] >> (2 * j)) & 0x03030303u) | ((((uint32_t)hm[4 + k] >> (4 * n + j)) & 0x01010101u) << 2);
                        s0 = dp4a_us(v0, a.q[k], s0);
                        s1 = dp4a_us(v1, a.q[4 + k], s1);
                    }
                    if constexpr (DBG) dbg[m * dms + 4 * n + j] = s0 + s1;
                    acc3[m] = fmaf(a.d8, i2f_small(s0 * scv[0] + s1 * scv[1]), acc3[m]);
                }
            }
        }
#pragma unroll
        for (int m = 0; m < M; ++m) acc[m] = fmaf(d, acc3[m], acc[m]);
    } else if constexpr (TYPE == T_Q4_K || TYPE == T_Q5_K) {
        constexpr int QS0 = (TYPE == T_Q4_K) ? 16 : 48;
        int hd[4];
        lds_words<4, A16>(wp, hd);
        const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&hd[0]));
        // 6-bit (scale, min) pairs, reference ggml-quants.c:625-632, four at a time as packed bytes
        const uint32_t h1 = (uint32_t)hd[1], h2 = (uint32_t)hd[2], h3 = (uint32_t)hd[3];
        const uint32_t sc4[2] = { h1 & 0x3F3F3F3Fu, (h3 & 0x0F0F0F0Fu) | ((h1 >> 2) & 0x30303030u) };
        const uint32_t mn4[2] = { h2 & 0x3F3F3F3Fu, ((h3 >> 4) & 0x0F0F0F0Fu) | ((h2 >> 2) & 0x30303030u) };
        int qh[8];
        if constexpr (TYPE == T_Q5_K) lds_words<8, A16>(wp + 16, qh);
        float accd[M], accm[M];
#pragma unroll
        for (int m = 0; m < M; ++m) { accd[m] = 0.0f; accm[m] = 0.0f; }
#pragma unroll
        for (int p = 0; p < 4; ++p) {
            int w[8];
            lds_words<8, A16>(wp + QS0 + 32 * p, w);
            uint32_t lo[8], hi[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if constexpr (TYPE == T_Q4_K) {
                    lo[j] = (uint32_t)w[j] & 0x0F0F0F0Fu;
                    hi[j] = (uint32_t)w[j] & 0xF0F0F0F0u;           // 16 x nibble, divided out after the dot
                } else {
                    const uint32_t b0 = (p < 2 ? ((uint32_t)qh[j] << (4 - 2 * p)) : ((uint32_t)qh[j] >> (2 * p - 4))) & 0x10101010u;
                    const uint32_t b1 = (p < 2 ? ((uint32_t)qh[j] << (3 - 2 * p)) : ((uint32_t)qh[j] >> (2 * p - 3))) & 0x10101010u;
                    lo[j] = ((uint32_t)w[j] & 0x0F0F0F0Fu) | b0;
                    hi[j] = (((uint32_t)w[j] >> 4) & 0x0F0F0F0Fu) | b1;
                }
            }
            const float sc0 = byte2f(sc4[p >> 1], 2 * (p & 1)), sc1 = byte2f(sc4[p >> 1], 2 * (p & 1) + 1);
            const float mn0 = byte2f(mn4[p >> 1], 2 * (p & 1)), mn1 = byte2f(mn4[p >> 1], 2 * (p & 1) + 1);
#pragma unroll
            for (int m = 0; m < M; ++m) {
                act_blk a0, a1;
                lds_act(ap + m * acs + (2 * p) * ACT_BLK, a0);
                lds_act(ap + m * acs + (2 * p + 1) * ACT_BLK, a1);
                int s0 = 0, s1 = 0;
#pragma unroll
                for (int j = 0; j < 8; ++j) { s0 = dp4a_us(lo[j], a0.q[j], s0); s1 = dp4a_us(hi[j], a1.q[j], s1); }
                if constexpr (TYPE == T_Q4_K) s1 >>= 4;
                if constexpr (DBG) { dbg[m * dms + 2 * p] = s0; dbg[m * dms + 2 * p + 1] = s1; }
                accd[m] = fmaf(a0.d8, i2f_small(s0) * sc0, accd[m]);
                accd[m] = fmaf(a1.d8, i2f_small(s1) * sc1, accd[m]);
                accm[m] = fmaf(a0.dsq, mn0, accm[m]);
                accm[m] = fmaf(a1.dsq, mn1, accm[m]);
            }
        }
#pragma unroll
        for (int m = 0; m < M; ++m) acc[m] += dm.x * accd[m] - dm.y * accm[m];
    } else {  // Q6_K: ql[128] qh[64] scales[16] d
        const float d = half_bits_to_float(lds16(wp + 208));
        float acc6[M];
#pragma unroll
        for (int m = 0; m < M; ++m) acc6[m] = 0.0f;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            int ql[16], qh[8], scw[2];
            lds_words_a2<16>(wp + 64 * h, ql);
            lds_words_a2<8>(wp + 128 + 32 * h, qh);
            lds_words_a2<2>(wp + 192 + 8 * h, scw);
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                uint32_t v[8];
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                    const uint32_t l4 = (uint32_t)ql[8 * (g & 1) + j], h4 = (uint32_t)qh[j];
                    const uint32_t nib = (g < 2 ? l4 : (l4 >> 4)) & 0x0F0F0F0Fu;
                    const uint32_t top = (g == 0 ? (h4 << 4) : g == 1 ? (h4 << 2) : g == 2 ? h4 : (h4 >> 2)) & 0x30303030u;
                    v[j] = nib | top;
                }
                const int sca = (int)(int8_t)((uint32_t)scw[g >> 1] >> (16 * (g & 1)));
                const int scb = (int)(int8_t)((uint32_t)scw[g >> 1] >> (16 * (g & 1) + 8));
#pragma unroll
                for (int m = 0; m < M; ++m) {
                    act_blk a; lds_act(ap + m * acs + (4 * h + g) * ACT_BLK, a);
                    int sa = a.m32a, sb = a.m32b;       // -32 * sum q: folds the -32 offset of the 6-bit code
#pragma unroll
                    for (int j = 0; j < 4; ++j) { sa = dp4a_us(v[j], a.q[j], sa); sb = dp4a_us(v[4 + j], a.q[4 + j], sb); }
                    if constexpr (DBG) dbg[m * dms + 4 * h + g] = sa + sb;
                    acc6[m] = fmaf(a.d8, (float)(sa * sca + sb * scb), acc6[m]);
                }
            }
        }
#pragma unroll
        for (int m = 0; m < M; ++m) acc[m] = fmaf(d, acc6[m], acc[m]);
    }
}

// ---- fused prologue: quantise M activation columns into shared memory ---------------------------
// q = roundf(x / d) with d = amax / 127, bit-identical to the reference (quantize.cu:38-47), at a
// fraction of the IEEE-division cost: r' = x * rcp(d) is within 3e-5 of x/d, so unless r' lies
// within 1e-4 of a rounding boundary (k + 0.5) its nearest integer IS roundf(x/d); the rare
// boundary cases (and denormal scales) take the exact division.
__device__ __forceinline__ int q8_round(float x, float d, float inv, bool exact_only) {
    const float r = x * inv;
    const float fm = r + 12582912.0f;              // 1.5 * 2^23: the mantissa now holds rint(r)
    const float u = r - (fm - 12582912.0f);
    if (exact_only || !(fabsf(u) <= 0.4999f)) return (int)roundf(__fdiv_rn(x, d));
    return __float_as_int(fm) - 0x4B400000;
}
// The same decision for the four elements of a lane with ONE branch: the fast roundings are straight-line code (sixteen of them
// interleave freely in quant_block_store_n), the exact division runs only when one of the four asks for it (about 1 lane in 1000).
__device__ __forceinline__ int q8_round4(const float4 x, float d, float inv, bool exact_only) {
    const float r0 = x.x * inv, r1 = x.y * inv, r2 = x.z * inv, r3 = x.w * inv;
    const float f0 = r0 + 12582912.0f, f1 = r1 + 12582912.0f, f2 = r2 + 12582912.0f, f3 = r3 + 12582912.0f;
    const float u0 = r0 - (f0 - 12582912.0f), u1 = r1 - (f1 - 12582912.0f), u2 = r2 - (f2 - 12582912.0f), u3 = r3 - (f3 - 12582912.0f);
    int q0 = __float_as_int(f0) - 0x4B400000, q1 = __float_as_int(f1) - 0x4B400000;
    int q2 = __float_as_int(f2) - 0x4B400000, q3 = __float_as_int(f3) - 0x4B400000;
    const bool ok = fabsf(u0) <= 0.4999f && fabsf(u1) <= 0.4999f && fabsf(u2) <= 0.4999f && fabsf(u3) <= 0.4999f;
    if (exact_only || !ok) {
        q0 = q8_round(x.x, d, inv, exact_only); q1 = q8_round(x.y, d, inv, exact_only);
        q2 = q8_round(x.z, d, inv, exact_only); q3 = q8_round(x.w, d, inv, exact_only);
    }
    return (int)__byte_perm(__byte_perm(q0, q1, 0x0040), __byte_perm(q2, q3, 0x0040), 0x5410);
}

// One 32-block held by 8 consecutive lanes (lane `sub` of the group has elements 4*sub..4*sub+3): quantise it with q8_1
// arithmetic and store it as block jb of the activation layout at `am` (see ACT_ITEM / ACT_BLK).  All 8 lanes must call.
__device__ __forceinline__ void quant_block_store(const float4 x, int jb, int sub, bool in_range, uint8_t * am, uint8_t * am2 = nullptr) {
    float amax = fmaxf(fmaxf(fabsf(x.x), fabsf(x.y)), fmaxf(fabsf(x.z), fabsf(x.w)));
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
    const float d = __fdiv_rn(amax, 127.0f);
    float inv; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(d));
    if (amax == 0.0f) inv = 0.0f;                                                  // an all-zero block: every q is 0, no branch
    const int word = q8_round4(x, d, inv, amax < 1e-20f && amax != 0.0f);
    int s = __dp4a(word, 0x01010101, 0);
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);          // sum of 16 elements (lanes 0-3 / 4-7)
    const int s16 = s;
    s += __shfl_xor_sync(0xffffffffu, s, 4);          // sum of the 32-block
    const int s16b = __shfl_down_sync(0xffffffffu, s16, 4);
    if (in_range) {
        const size_t off = (size_t)(jb >> 3) * ACT_ITEM + (jb & 7) * ACT_BLK;
        reinterpret_cast<int *>(am + off)[sub] = word;
        if (am2) reinterpret_cast<int *>(am2 + off)[sub] = word;
        if (sub == 0) {
            const float d8 = __half2float(__float2half_rn(d));   // q8_1 stores d as binary16 (quantize.cu:47)
            int4 t; t.x = __float_as_int(d8); t.y = __float_as_int(d8 * (float)s); t.z = -32 * s16; t.w = -32 * s16b;
            *reinterpret_cast<int4 *>(am + off + 32) = t;
            if (am2) *reinterpret_cast<int4 *>(am2 + off + 32) = t;
        }
    }
}

// SILU as every kernel that fuses it computes it (unary_mul_kernel, the gate / up epilogue of the batch-1 mat-vec): one expf, one
// IEEE division, nothing the compiler could contract differently from one kernel to the next
__device__ __forceinline__ float silu_fn(float x) { return x / (1.0f + expf(-x)); }

// Straight-line part of q8_round4: the word of fast roundings and whether all four are trustworthy (no branch in here).
__device__ __forceinline__ int q8_fast4(const float4 x, float inv, bool & ok) {
    const float r0 = x.x * inv, r1 = x.y * inv, r2 = x.z * inv, r3 = x.w * inv;
    const float f0 = r0 + 12582912.0f, f1 = r1 + 12582912.0f, f2 = r2 + 12582912.0f, f3 = r3 + 12582912.0f;
    const float u0 = r0 - (f0 - 12582912.0f), u1 = r1 - (f1 - 12582912.0f), u2 = r2 - (f2 - 12582912.0f), u3 = r3 - (f3 - 12582912.0f);
    ok = ok && fabsf(u0) <= 0.4999f && fabsf(u1) <= 0.4999f && fabsf(u2) <= 0.4999f && fabsf(u3) <= 0.4999f;
    const int q0 = __float_as_int(f0) - 0x4B400000, q1 = __float_as_int(f1) - 0x4B400000;
    const int q2 = __float_as_int(f2) - 0x4B400000, q3 = __float_as_int(f3) - 0x4B400000;
    return (int)__byte_perm(__byte_perm(q0, q1, 0x0040), __byte_perm(q2, q3, 0x0040), 0x5410);
}
static __device__ __noinline__ int q8_exact4(const float4 x, float d, float inv, bool tiny) {      // the rare path, kept out of line
    const int q0 = q8_round(x.x, d, inv, tiny), q1 = q8_round(x.y, d, inv, tiny), q2 = q8_round(x.z, d, inv, tiny), q3 = q8_round(x.w, d, inv, tiny);
    return (int)__byte_perm(__byte_perm(q0, q1, 0x0040), __byte_perm(q2, q3, 0x0040), 0x5410);
}

// NP 32-blocks, each held by FOUR consecutive lanes with eight consecutive elements per lane (a = 8*sub4 .. +3, b = 8*sub4+4 .. +7):
// the same q8_1 arithmetic as quant_block_store (max, integer sums and per-element roundings do not depend on which lane holds
// what), with 2.4x fewer instructions per element (two butterfly rounds instead of three, one division / reciprocal per eight
// elements) and ONE branch for all the roundings, so the NP dependent chains overlap.  Used where every CTA of a launch repeats
// the quantisation (the norm prologue of the batch-1 GEMV), see profiles/r02_norm_prologue.md.  in[n] uniform over the 4 lanes.
template <int NP>
__device__ __forceinline__ void quant_block_store8(const float4 (&a)[NP], const float4 (&b)[NP], const int (&jb)[NP], int sub4, const bool (&in)[NP],
                                                   uint8_t * am, uint8_t * am2 = nullptr) {
    float amax[NP], d[NP], inv[NP];
    int w0[NP], w1[NP], s[NP], s16[NP], s16b[NP];
    bool ok = true;
#pragma unroll
    for (int n = 0; n < NP; ++n)
        amax[n] = fmaxf(fmaxf(fmaxf(fabsf(a[n].x), fabsf(a[n].y)), fmaxf(fabsf(a[n].z), fabsf(a[n].w))),
                        fmaxf(fmaxf(fabsf(b[n].x), fabsf(b[n].y)), fmaxf(fabsf(b[n].z), fabsf(b[n].w))));
#pragma unroll
    for (int n = 0; n < NP; ++n) amax[n] = fmaxf(amax[n], __shfl_xor_sync(0xffffffffu, amax[n], 1));
#pragma unroll
    for (int n = 0; n < NP; ++n) amax[n] = fmaxf(amax[n], __shfl_xor_sync(0xffffffffu, amax[n], 2));
#pragma unroll
    for (int n = 0; n < NP; ++n) {
        d[n] = __fdiv_rn(amax[n], 127.0f);
        asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv[n]) : "f"(d[n]));
        if (amax[n] == 0.0f) inv[n] = 0.0f;                            // an all-zero block: every q is 0
        ok = ok && !(amax[n] < 1e-20f && amax[n] != 0.0f);             // denormal scales take the exact division
    }
#pragma unroll
    for (int n = 0; n < NP; ++n) { w0[n] = q8_fast4(a[n], inv[n], ok); w1[n] = q8_fast4(b[n], inv[n], ok); }
    if (!ok) {
#pragma unroll
        for (int n = 0; n < NP; ++n) {
            const bool tiny = amax[n] < 1e-20f && amax[n] != 0.0f;
            w0[n] = q8_exact4(a[n], d[n], inv[n], tiny); w1[n] = q8_exact4(b[n], d[n], inv[n], tiny);
        }
    }
#pragma unroll
    for (int n = 0; n < NP; ++n) s[n] = __dp4a(w1[n], 0x01010101, __dp4a(w0[n], 0x01010101, 0));
#pragma unroll
    for (int n = 0; n < NP; ++n) { s[n] += __shfl_xor_sync(0xffffffffu, s[n], 1); s16[n] = s[n]; }      // 16 elements (lanes 0-1 / 2-3)
#pragma unroll
    for (int n = 0; n < NP; ++n) { s[n] += __shfl_xor_sync(0xffffffffu, s[n], 2); s16b[n] = __shfl_down_sync(0xffffffffu, s16[n], 2); }
#pragma unroll
    for (int n = 0; n < NP; ++n) {
        if (in[n]) {
            const size_t off = (size_t)(jb[n] >> 3) * ACT_ITEM + (jb[n] & 7) * ACT_BLK;
            reinterpret_cast<int2 *>(am + off)[sub4] = make_int2(w0[n], w1[n]);
            if (am2) reinterpret_cast<int2 *>(am2 + off)[sub4] = make_int2(w0[n], w1[n]);
            if (sub4 == 0) {
                const float d8 = __half2float(__float2half_rn(d[n]));
                int4 t; t.x = __float_as_int(d8); t.y = __float_as_int(d8 * (float)s[n]); t.z = -32 * s16[n]; t.w = -32 * s16b[n];
                *reinterpret_cast<int4 *>(am + off + 32) = t;
                if (am2) *reinterpret_cast<int4 *>(am2 + off + 32) = t;
            }
        }
    }
}

// NB independent 32-blocks per 8-lane group, the same arithmetic per block as quant_block_store, written so that the NB dependent
// chains (3 + 4 shuffles, an IEEE division, a reciprocal, four roundings: ~1500 cycles of latency each) advance in lockstep instead
// of one after the other.  The fused norm prologue of the batch-1 GEMV runs 4 blocks per thread with 2 warps per scheduler: called
// four times in a row the quantiser alone took 3.6 us of the prologue's 6 (timeline of the chain kernel, profiles/r02_chain_timeline.log).
// in[n] must be uniform over the 8 lanes of a group; all lanes of the warp must call.
template <int NB>
__device__ __forceinline__ void quant_block_store_n(const float4 (&x)[NB], const int (&jb)[NB], int sub, const bool (&in)[NB], uint8_t * am, uint8_t * am2 = nullptr) {
    float amax[NB], d[NB], inv[NB];
    int word[NB], s[NB], s16[NB], s16b[NB];
#pragma unroll
    for (int n = 0; n < NB; ++n) amax[n] = fmaxf(fmaxf(fabsf(x[n].x), fabsf(x[n].y)), fmaxf(fabsf(x[n].z), fabsf(x[n].w)));
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
#pragma unroll
        for (int n = 0; n < NB; ++n) amax[n] = fmaxf(amax[n], __shfl_xor_sync(0xffffffffu, amax[n], o));
    }
#pragma unroll
    for (int n = 0; n < NB; ++n) d[n] = __fdiv_rn(amax[n], 127.0f);
#pragma unroll
    for (int n = 0; n < NB; ++n) asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(inv[n]) : "f"(d[n]));
#pragma unroll
    for (int n = 0; n < NB; ++n) {
        if (amax[n] == 0.0f) inv[n] = 0.0f;
        word[n] = q8_round4(x[n], d[n], inv[n], amax[n] < 1e-20f && amax[n] != 0.0f);
        s[n] = __dp4a(word[n], 0x01010101, 0);
    }
#pragma unroll
    for (int n = 0; n < NB; ++n) s[n] += __shfl_xor_sync(0xffffffffu, s[n], 1);
#pragma unroll
    for (int n = 0; n < NB; ++n) s[n] += __shfl_xor_sync(0xffffffffu, s[n], 2);
#pragma unroll
    for (int n = 0; n < NB; ++n) { s16[n] = s[n]; s[n] += __shfl_xor_sync(0xffffffffu, s[n], 4); }
#pragma unroll
    for (int n = 0; n < NB; ++n) s16b[n] = __shfl_down_sync(0xffffffffu, s16[n], 4);
#pragma unroll
    for (int n = 0; n < NB; ++n) {
        if (in[n]) {
            const size_t off = (size_t)(jb[n] >> 3) * ACT_ITEM + (jb[n] & 7) * ACT_BLK;
            reinterpret_cast<int *>(am + off)[sub] = word[n];
            if (am2) reinterpret_cast<int *>(am2 + off)[sub] = word[n];
            if (sub == 0) {
                const float d8 = __half2float(__float2half_rn(d[n]));
                int4 t; t.x = __float_as_int(d8); t.y = __float_as_int(d8 * (float)s[n]); t.z = -32 * s16[n]; t.w = -32 * s16b[n];
                *reinterpret_cast<int4 *>(am + off + 32) = t;
                if (am2) *reinterpret_cast<int4 *>(am2 + off + 32) = t;
            }
        }
    }
}

// Eight lanes per 32-block (a float4 each); nthreads threads (a multiple of 32) cooperate.  Must be
// called by all of them.  nb32 = K/32 blocks per column.
__device__ __forceinline__ void quantize_acts(const float * __restrict__ X, int64_t xs, int M, int nb32, bool vec_ok,
                                              uint8_t * act, uint32_t acs, int tid, int nthreads) {
    const int sub = tid & 7, grp = tid >> 3, ngrp = nthreads >> 3;
    constexpr int UN = 4;                                    // independent loads in flight per lane
    for (int m = 0; m < M; ++m) {
        const float * xm = X + (int64_t)m * xs + 4 * sub;
        uint8_t * am = act + (size_t)m * acs;
        for (int base = 0; base < nb32; base += UN * ngrp) {
            float4 xv[UN];
#pragma unroll
            for (int u = 0; u < UN; ++u) {
                const int jb = base + u * ngrp + grp;
                xv[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                if (jb < nb32) {
                    const float * src = xm + 32 * jb;
                    if (vec_ok) xv[u] = __ldg(reinterpret_cast<const float4 *>(src));
                    else { xv[u].x = __ldg(src); xv[u].y = __ldg(src + 1); xv[u].z = __ldg(src + 2); xv[u].w = __ldg(src + 3); }
                }
            }
#pragma unroll
            for (int u = 0; u < UN; ++u) {
                const int jb = base + u * ngrp + grp;         // uniform across the 8 lanes of a group
                quant_block_store(xv[u], jb, sub, jb < nb32, am);
            }
        }
    }
}

}  // namespace b200
