// dequant.cuh -- element-wise dequantisation of GGUF blocks, bit-exact to the reference's
// dequantize_row_* (ggml/src/ggml-quants.c:249-267, 269-341 (q4_1 q5_0 q5_1), 343-357, 1274-1296, 1480-1503, 1684-1713).
#pragma once
#include "common.cuh"

namespace b200 {

// 6-bit scale/min unpack (reference ggml-quants.c:625-632)
__device__ __forceinline__ void scale_min_k4(int j, const uint8_t * q, int & sc, int & mn) {
    if (j < 4) { sc = q[j] & 63; mn = q[j + 4] & 63; }
    else { sc = (q[j + 4] & 0x0F) | ((q[j - 4] >> 6) << 4); mn = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
}

__device__ __forceinline__ float ldh(const uint8_t * p) {   // 2-byte aligned half load
    return __half2float(__ushort_as_half(*reinterpret_cast<const unsigned short *>(p)));
}

// the table of the 4-bit non-linear formats (reference ggml-common.h:1077-1079)
__device__ const int8_t iq4_values[16] = { -127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113 };

// One element per thread; every product/difference is an explicit single rounding so the result is
// bit-identical to the reference's unfused CPU arithmetic (ggml-quants.c dequantize_row_*).
template <int TYPE> __device__ __forceinline__ float dequant_elem(const uint8_t * row, int64_t e) {
    if constexpr (TYPE == T_Q4_0) {
        const uint8_t * b = row + (e >> 5) * 18; const int j = e & 31;
        const int q = j < 16 ? (b[2 + j] & 0xF) : (b[2 + j - 16] >> 4);
        return __fmul_rn((float)(q - 8), ldh(b));
    } else if constexpr (TYPE == T_Q4_1) {
        const uint8_t * b = row + (e >> 5) * 20; const int j = e & 31;
        const int q = j < 16 ? (b[4 + j] & 0xF) : (b[4 + j - 16] >> 4);
        return __fadd_rn(__fmul_rn((float)q, ldh(b)), ldh(b + 2));
    } else if constexpr (TYPE == T_Q5_0 || TYPE == T_Q5_1) {
        constexpr int BB = TYPE == T_Q5_0 ? 22 : 24, QH = TYPE == T_Q5_0 ? 2 : 4;
        const uint8_t * b = row + (e >> 5) * BB; const int j = e & 31;
        const int nib = j < 16 ? (b[QH + 4 + j] & 0xF) : (b[QH + 4 + j - 16] >> 4);
        const int q = nib | (((b[QH + (j >> 3)] >> (j & 7)) & 1) << 4);          // bit j of the little-endian qh word
        if constexpr (TYPE == T_Q5_0) return __fmul_rn((float)(q - 16), ldh(b));
        else return __fadd_rn(__fmul_rn((float)q, ldh(b)), ldh(b + 2));
    } else if constexpr (TYPE == T_IQ4_NL) {       // ggml-quants.c:2428-2444
        const uint8_t * b = row + (e >> 5) * 18; const int j = e & 31;
        const int idx = j < 16 ? (b[2 + j] & 0xF) : (b[2 + j - 16] >> 4);
        return __fmul_rn(ldh(b), (float)iq4_values[idx]);
    } else if constexpr (TYPE == T_IQ4_XS) {       // ggml-quants.c:2446-2467
        const uint8_t * b = row + (e >> 8) * 136; const int r = e & 255, ib = r >> 5, j = r & 31;
        const unsigned sh = b[2] | (b[3] << 8);
        const int ls = (int)(((b[4 + ib / 2] >> (4 * (ib % 2))) & 0xF) | (((sh >> (2 * ib)) & 3) << 4)) - 32;
        const int idx = j < 16 ? (b[8 + 16 * ib + j] & 0xF) : (b[8 + 16 * ib + j - 16] >> 4);
        return __fmul_rn(__fmul_rn(ldh(b), (float)ls), (float)iq4_values[idx]);
    } else if constexpr (TYPE == T_Q2_K) {         // ggml-quants.c:706-736
        const uint8_t * b = row + (e >> 8) * 84; const int r = e & 255, n = r >> 7, j = (r >> 5) & 3, l = r & 31;
        const uint8_t sc = b[8 * n + 2 * j + (l >> 4)];
        const float dl = __fmul_rn(ldh(b + 80), (float)(sc & 0xF)), ml = __fmul_rn(ldh(b + 82), (float)(sc >> 4));
        return __fsub_rn(__fmul_rn(dl, (float)((b[16 + 32 * n + l] >> (2 * j)) & 3)), ml);
    } else if constexpr (TYPE == T_Q3_K) {         // ggml-quants.c:1050-1098
        const uint8_t * b = row + (e >> 8) * 110; const int r = e & 255, n = r >> 7, j = (r >> 5) & 3, l = r & 31;
        const int is = r >> 4, sb = is & 3, sg = is >> 2;
        const int lo = sg < 2 ? (b[96 + 4 * sg + sb] & 0xF) : (b[96 + 4 * (sg - 2) + sb] >> 4);
        const int sc = (lo | (((b[104 + sb] >> (2 * sg)) & 3) << 4)) - 32;
        const int q = (int)((b[32 + 32 * n + l] >> (2 * j)) & 3) - ((b[l] & (1 << (4 * n + j))) ? 0 : 4);
        return __fmul_rn(__fmul_rn(ldh(b + 108), (float)sc), (float)q);
    } else if constexpr (TYPE == T_Q8_0) {
        const uint8_t * b = row + (e >> 5) * 34;
        return __fmul_rn((float)(int8_t)b[2 + (e & 31)], ldh(b));
    } else if constexpr (TYPE == T_Q4_K || TYPE == T_Q5_K) {
        constexpr int BB = TYPE == T_Q4_K ? 144 : 176, QS = TYPE == T_Q4_K ? 16 : 48;
        const uint8_t * b = row + (e >> 8) * BB; const int r = e & 255, s = r >> 5, l = r & 31, p = s >> 1;
        int sc, mn; scale_min_k4(s, b + 4, sc, mn);
        const uint8_t qb = b[QS + 32 * p + l];
        int q = (s & 1) ? (qb >> 4) : (qb & 0xF);
        if constexpr (TYPE == T_Q5_K) q += ((b[16 + l] >> s) & 1) ? 16 : 0;
        const float d = __fmul_rn(ldh(b), (float)sc), m = __fmul_rn(ldh(b + 2), (float)mn);
        return __fsub_rn(__fmul_rn(d, (float)q), m);
    } else {  // Q6_K
        const uint8_t * b = row + (e >> 8) * 210; const int r0 = e & 255, h = r0 >> 7, r = r0 & 127, g = r >> 5, l = r & 31;
        const uint8_t lo = b[64 * h + 32 * (g & 1) + l];
        const int nib = g < 2 ? (lo & 0xF) : (lo >> 4);
        const int hi = (b[128 + 32 * h + l] >> (2 * g)) & 3;
        const int q = (int)(int8_t)(nib | (hi << 4)) - 32;
        const float ds = __fmul_rn(ldh(b + 208), (float)(int8_t)b[192 + (r0 >> 4)]);
        return __fmul_rn(ds, (float)q);
    }
}

}  // namespace b200
