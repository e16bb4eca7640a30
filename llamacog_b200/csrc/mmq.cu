// mmq.cu -- batched quantized matmul for prefill (reference ggml-cuda/mmq.cu:1-366,
// mmq.cuh tile loaders + mma.sync; reference dispatch ggml-cuda.cu:1936-2005) rebuilt for sm_100a:
//
//   dst[m][n] = sum_k dequant(W[n][k]) * X[m][k]          W: GGUF blocks, X: f32 activations, dst: f32
//
// One CTA owns a 128-weight-row x TN-token output tile whose f32 accumulator lives in TMEM.  Per 256-K
// block:  warp 0 TMA-loads the raw GGUF bytes of the 128 rows (2-D tensor map over the weight bytes),
// warp 3 TMA-loads the bf16 activation tiles (2-D tensor map, 128-byte swizzle), eight dequantiser warps
// expand the blocks to bf16 and store them with tcgen05.st into a ring of A-operand slots in TENSOR
// MEMORY (thread = weight row = TMEM lane), and one thread of warp 1 issues tcgen05.mma with A from TMEM
// and B from shared memory (kind::f16, bf16 x bf16 -> f32, M=128, N=TN, K=16).  Keeping A out of shared
// memory matters: an SS-mode MMA reads A and B from shared memory at 96 B/clk of the SM's 128 B/clk, which
// left the first version of this kernel shared-memory bound at 37 % tensor-pipe utilisation
// (profiles/r01_mmq_notes.md).  Raw bytes are read from HBM exactly once per tile row.
// Activations are converted f32 -> bf16 once per mat-mul by a small pre-pass (the reference quantises
// them to q8_1 there instead, ggml-cuda/quantize.cu).
#include "tc.cuh"
#include "../../include/b200_ops.h"

#include <cuda.h>
#include <cuda_bf16.h>
#include <cstdlib>
#include <cstring>
#include <type_traits>

namespace b200 {

__device__ __forceinline__ uint32_t fma_bf16x2(uint32_t a, uint32_t b, uint32_t c) {
    uint32_t d; asm("fma.rn.bf16x2 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c)); return d;
}
// {f, f} as packed bf16 (round to nearest even)
__device__ __forceinline__ uint32_t pack_bf16_dup(float f) {
    uint32_t d; asm("cvt.rn.bf16x2.f32 %0, %1, %1;" : "=r"(d) : "f"(f)); return d;
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
    uint32_t d; asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo)); return d;
}
__device__ __forceinline__ float bf16lo_to_float(uint32_t packed) { return __uint_as_float(packed << 16); }

// ------------------------------------------------------------------------------------------------
// Dequantisers.  A dequantiser thread owns one weight row r of the tile and one half h of each 256-K
// block (h = 0: K 0..127, h = 1: K 128..255), i.e. the two 64-K sub-chunks c = 2h, 2h+1.  For each it
// produces the row's 64 bf16 as 32 packed registers (chunk u = elements 8u..8u+7 = out[4u..4u+3]), which
// one tcgen05.st writes to the 32 columns of the thread's TMEM lane.
//
// Integer codes become bf16 without conversions: a byte 0x80|(q << s) under a 0x3F high byte is the
// bf16 value 1 + q/2^n (n = 7-s).  Subtracting 1.5 is exact and centres the code, (q - 2^(n-1))/2^n, so
//   w = q*scale - offset = ((1 + q/2^n) - 1.5) * S + C,   S = 2^n*scale,  C = S/2 - offset
// costs one packed add and one packed fma per two weights; formats without an offset (Q6_K, Q4_0: the
// centring IS their zero point) need only a multiply.  S and C are rounded to bf16 once per sub-block and
// C is computed from the rounded S.  (Values follow dequantize_row_*, reference ggml-quants.c; the result
// is rounded to bf16 -- the op's parity bar is NMSE <= 5e-4, tests/test-backend-ops.cpp.)
// ------------------------------------------------------------------------------------------------
#define PRMT_01 0x4140u   // {b0, 0x3F, b1, 0x3F}
#define PRMT_23 0x4342u   // {b2, 0x3F, b3, 0x3F}

// four code bytes (already shifted, 0x80 set) -> two packed bf16 pairs:  ((1 + q/2^n) - 1.5) * S + C.  The subtraction is exact
// (it centres the code: (q - 2^(n-1)) / 2^n), so the only roundings are those of S, C and the result; C = S/2 - offset is small
// for the roughly symmetric blocks a quantiser produces, which keeps its rounding error far below one code step.
__device__ __forceinline__ void expand4(uint32_t codes, uint32_t s2, uint32_t c2, uint32_t & o01, uint32_t & o23) {
    constexpr uint32_t M15 = 0xBFC0BFC0u;      // {-1.5, -1.5} bf16
    uint32_t a, b;
    asm("add.rn.bf16x2 %0, %1, %2;" : "=r"(a) : "r"(__byte_perm(codes, 0x3F3F3F3Fu, PRMT_01)), "r"(M15));
    asm("add.rn.bf16x2 %0, %1, %2;" : "=r"(b) : "r"(__byte_perm(codes, 0x3F3F3F3Fu, PRMT_23)), "r"(M15));
    o01 = fma_bf16x2(a, s2, c2);
    o23 = fma_bf16x2(b, s2, c2);
}
// symmetric formats (no per-block minimum): remove the code offset exactly first ((1 + q/2^n) - off is exact in bf16),
// then one multiply -- a single rounding per weight
__device__ __forceinline__ uint32_t add_bf16x2(uint32_t a, uint32_t b) {
    uint32_t d; asm("add.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d;
}
__device__ __forceinline__ uint32_t mul_bf16x2(uint32_t a, uint32_t b) {
    uint32_t d; asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(a), "r"(b)); return d;
}
__device__ __forceinline__ void expand4_sym(uint32_t codes, uint32_t s2, uint32_t & o01, uint32_t & o23) {
    constexpr uint32_t M15 = 0xBFC0BFC0u;      // {-1.5, -1.5} bf16
    o01 = mul_bf16x2(add_bf16x2(__byte_perm(codes, 0x3F3F3F3Fu, PRMT_01), M15), s2);
    o23 = mul_bf16x2(add_bf16x2(__byte_perm(codes, 0x3F3F3F3Fu, PRMT_23), M15), s2);
}
// scale pair from S (float): s2 = {S,S} bf16, c2 = {Sr/2 - off} with Sr the rounded S
__device__ __forceinline__ void make_sb(float S, float off, uint32_t & s2, uint32_t & c2) {
    s2 = pack_bf16_dup(S);
    c2 = pack_bf16_dup(fmaf(bf16lo_to_float(s2), 0.5f, -off));
}
__device__ __forceinline__ float ubyte_f(uint32_t w, int i) { return (float)((w >> (8 * i)) & 0xFFu); }

template <int TYPE> struct dq;

// ---- Q4_K: d, dmin (f16) | scales[12] (6-bit sc/mn x 8) | qs[128]; sub-block j (32 elems): nibble (j&1) of qs[32*(j/2) + l]
template <> struct dq<T_Q4_K> {
    static constexpr int PITCH = 144;      // bytes of one row's block in the raw stage
    static constexpr int STEP16 = 72;      // block stride in 16-bit elements (tensor-map coordinate)
    static constexpr bool SHIFTED = false;
    uint32_t s2[4], b2[4];
    template <int S> __device__ __forceinline__ void prep(uint32_t row, int h) {
        const int4 hd = lds128(row);
        const float d = half_bits_to_float(hd.x & 0xFFFF), dmin = half_bits_to_float((uint32_t)hd.x >> 16);
        uint32_t SC, MN;
        if (h == 0) { SC = hd.y & 0x3F3F3F3Fu; MN = hd.z & 0x3F3F3F3Fu; }
        else { SC = (hd.w & 0x0F0F0F0Fu) | (((uint32_t)hd.y >> 2) & 0x30303030u); MN = (((uint32_t)hd.w >> 4) & 0x0F0F0F0Fu) | (((uint32_t)hd.z >> 2) & 0x30303030u); }
#pragma unroll
        for (int j = 0; j < 4; ++j) make_sb(16.0f * d * ubyte_f(SC, j), dmin * ubyte_f(MN, j), s2[j], b2[j]);
    }
    __device__ __forceinline__ void emit(uint32_t row, int h, int t, uint32_t (&out)[32]) const {
        const uint32_t qs = row + 16 + 32 * (2 * h + t);
        const int4 q0 = lds128(qs), q1 = lds128(qs + 16);
        const uint32_t w[8] = {(uint32_t)q0.x, (uint32_t)q0.y, (uint32_t)q0.z, (uint32_t)q0.w, (uint32_t)q1.x, (uint32_t)q1.y, (uint32_t)q1.z, (uint32_t)q1.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {      // low nibbles -> chunks 0..3 (sub-block 2t), high -> 4..7 (sub-block 2t+1)
            uint32_t o[4], p[4];
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const uint32_t x = w[2 * u + i];
                expand4(((x << 3) & 0x78787878u) | 0x80808080u, s2[2 * t], b2[2 * t], o[2 * i], o[2 * i + 1]);
                expand4(((x >> 1) & 0x78787878u) | 0x80808080u, s2[2 * t + 1], b2[2 * t + 1], p[2 * i], p[2 * i + 1]);
            }
            { out[4 * (u)] = o[0]; out[4 * (u) + 1] = o[1]; out[4 * (u) + 2] = o[2]; out[4 * (u) + 3] = o[3]; }
            { out[4 * ((u + 4))] = p[0]; out[4 * ((u + 4)) + 1] = p[1]; out[4 * ((u + 4)) + 2] = p[2]; out[4 * ((u + 4)) + 3] = p[3]; }
        }
    }
};

// ---- Q5_K: d, dmin | scales[12] | qh[32] | qs[128]; as Q4_K plus bit j of qh[l] as the code's fifth bit
template <> struct dq<T_Q5_K> {
    static constexpr int PITCH = 176;
    static constexpr int STEP16 = 88;
    static constexpr bool SHIFTED = false;
    uint32_t s2[4], b2[4], qh[8];
    template <int S> __device__ __forceinline__ void prep(uint32_t row, int h) {
        const int4 hd = lds128(row);
        const float d = half_bits_to_float(hd.x & 0xFFFF), dmin = half_bits_to_float((uint32_t)hd.x >> 16);
        uint32_t SC, MN;
        if (h == 0) { SC = hd.y & 0x3F3F3F3Fu; MN = hd.z & 0x3F3F3F3Fu; }
        else { SC = (hd.w & 0x0F0F0F0Fu) | (((uint32_t)hd.y >> 2) & 0x30303030u); MN = (((uint32_t)hd.w >> 4) & 0x0F0F0F0Fu) | (((uint32_t)hd.z >> 2) & 0x30303030u); }
#pragma unroll
        for (int j = 0; j < 4; ++j) make_sb(32.0f * d * ubyte_f(SC, j), dmin * ubyte_f(MN, j), s2[j], b2[j]);
        const int4 a = lds128(row + 16), b = lds128(row + 32);
        const int sh = 4 * h;              // this half's four sub-block bits move to bits 0..3
        qh[0] = (uint32_t)a.x >> sh; qh[1] = (uint32_t)a.y >> sh; qh[2] = (uint32_t)a.z >> sh; qh[3] = (uint32_t)a.w >> sh;
        qh[4] = (uint32_t)b.x >> sh; qh[5] = (uint32_t)b.y >> sh; qh[6] = (uint32_t)b.z >> sh; qh[7] = (uint32_t)b.w >> sh;
    }
    __device__ __forceinline__ void emit(uint32_t row, int h, int t, uint32_t (&out)[32]) const {
        const uint32_t qs = row + 48 + 32 * (2 * h + t);
        const int4 q0 = lds128(qs), q1 = lds128(qs + 16);
        const uint32_t w[8] = {(uint32_t)q0.x, (uint32_t)q0.y, (uint32_t)q0.z, (uint32_t)q0.w, (uint32_t)q1.x, (uint32_t)q1.y, (uint32_t)q1.z, (uint32_t)q1.w};
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            uint32_t o[4], p[4];
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const uint32_t x = w[2 * u + i], hb = qh[2 * u + i];
                // code byte = 0x80 | q5 << 2 : nibble at bits 2..5, fifth bit at bit 6
                const uint32_t lo = ((x << 2) & 0x3C3C3C3Cu) | 0x80808080u | ((hb << (6 - 2 * t)) & 0x40404040u);
                const uint32_t hi = ((x >> 2) & 0x3C3C3C3Cu) | 0x80808080u | ((hb << (5 - 2 * t)) & 0x40404040u);
                expand4(lo, s2[2 * t], b2[2 * t], o[2 * i], o[2 * i + 1]);
                expand4(hi, s2[2 * t + 1], b2[2 * t + 1], p[2 * i], p[2 * i + 1]);
            }
            { out[4 * (u)] = o[0]; out[4 * (u) + 1] = o[1]; out[4 * (u) + 2] = o[2]; out[4 * (u) + 3] = o[3]; }
            { out[4 * ((u + 4))] = p[0]; out[4 * ((u + 4)) + 1] = p[1]; out[4 * ((u + 4)) + 2] = p[2]; out[4 * ((u + 4)) + 3] = p[3]; }
        }
    }
};

// NW 32-bit words at byte offset S (even, < 16, compile time) past a 16-byte aligned shared address
template <int NW, int S> __device__ __forceinline__ void lds_shifted(uint32_t a16, uint32_t (&out)[NW]) {
    constexpr int NV = (S + 4 * NW + 15) / 16, WO = S >> 2, SH = (S & 2) * 8;
    uint32_t w[4 * NV + 1];
#pragma unroll
    for (int v = 0; v < NV; ++v) {
        const int4 t = lds128(a16 + 16 * v);
        w[4 * v] = (uint32_t)t.x; w[4 * v + 1] = (uint32_t)t.y; w[4 * v + 2] = (uint32_t)t.z; w[4 * v + 3] = (uint32_t)t.w;
    }
    w[4 * NV] = 0;
#pragma unroll
    for (int i = 0; i < NW; ++i) out[i] = SH ? __funnelshift_r(w[WO + i], w[WO + i + 1], SH) : w[WO + i];
}

// ---- Q6_K: ql[128] | qh[64] | scales int8[16] | d (f16).  Element e: half hh = e/128, g = (e%128)/32, l = e%32:
//      nibble (g/2) of ql[64hh + 32(g&1) + l], bits 2g..2g+1 of qh[32hh + l]; value (q6 - 32) * d * scales[e/16].
//      A 210-byte block is not a multiple of 16: TMA fetches the enclosing 16-byte aligned 240-byte window, so block
//      kb starts S = (210 kb) mod 16 = 2 (kb mod 8) bytes into its row of the raw stage (compile-time per unrolled kb).
template <> struct dq<T_Q6_K> {
    static constexpr int BB = 210;        // bytes per block: block kb of a row starts (BB * kb) % 16 bytes into its 16-byte aligned window
    static constexpr int PITCH = 240;
    static constexpr int STEP16 = 105;
    static constexpr bool SHIFTED = true;
    uint32_t s2[8];                        // 8 x 16-element scale groups of this half
    uint32_t ql[16], qh[8];                // this half's 64 low-nibble bytes and 32 high-bit bytes
    template <int S> __device__ __forceinline__ void prep(uint32_t row, int h) {
        const float d = half_bits_to_float(lds16(row + 208 + S));
        uint32_t sc[4]; lds_shifted<4, S>(row + 192, sc);
        const uint32_t sx = h ? sc[2] : sc[0], sy = h ? sc[3] : sc[1];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int v = (int)(int8_t)(((j < 4 ? sx : sy) >> (8 * (j & 3))) & 0xFF);
            s2[j] = pack_bf16_dup(64.0f * d * (float)v);                  // (q6 - 32) * ds = ((1 + q6/64) - 1.5) * 64ds
        }
        lds_shifted<16, S>(row + 64 * h, ql);
        lds_shifted<8, S>(row + 128 + 32 * h, qh);
    }
    __device__ __forceinline__ void emit(uint32_t, int, int t, uint32_t (&out)[32]) const {
#pragma unroll
        for (int g = 0; g < 2; ++g) {      // 32 elements each; sub-chunk c = 2h + t holds groups 2t, 2t+1 of half h
            const int gg = 2 * t + g;      // bit pair 2gg of qh
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                uint32_t o[4];
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const uint32_t x = ql[8 * g + 2 * u + i], y = qh[2 * u + i];
                    // code byte = 0x80 | q6 << 1 : nibble at bits 1..4, high pair at bits 5..6
                    const uint32_t nib = t == 0 ? (x << 1) : (x >> 3);
                    const uint32_t hi2 = 2 * gg <= 5 ? (y << (5 - 2 * gg)) : (y >> (2 * gg - 5));
                    const uint32_t code = (nib & 0x1E1E1E1Eu) | 0x80808080u | (hi2 & 0x60606060u);
                    const int si = 4 * t + 2 * g + (u >> 1);
                    expand4_sym(code, s2[si], o[2 * i], o[2 * i + 1]);
                }
                { out[4 * ((4 * g + u))] = o[0]; out[4 * ((4 * g + u)) + 1] = o[1]; out[4 * ((4 * g + u)) + 2] = o[2]; out[4 * ((4 * g + u)) + 3] = o[3]; }
            }
        }
    }
};

// ---- Q2_K (round 2): scales[16] (4-bit scale | 4-bit min per 16 elements) | qs[64] | d | dmin; element 128n + 32j + l: bits 2j..2j+1 of
//      qs[32n + l]; q*d*sc - dmin*mn = ((1 + q/4) - 1.5) * 4 d sc + (2 d sc - dmin mn).  84-byte blocks: window of 96 bytes, shift (84 kb) % 16.
template <> struct dq<T_Q2_K> {
    static constexpr int BB = 84;
    static constexpr int PITCH = 96;
    static constexpr int STEP16 = 42;
    static constexpr bool SHIFTED = true;
    uint32_t s2[8], b2[8], qs[8];          // the eight 16-element scale groups and the 32 code bytes of this half
    template <int S> __device__ __forceinline__ void prep(uint32_t row, int h) {
        uint32_t dmw[1]; lds_shifted<1, S>(row + 80, dmw);
        const float d = half_bits_to_float(dmw[0] & 0xFFFFu), dmin = half_bits_to_float(dmw[0] >> 16);
        uint32_t sc[4]; lds_shifted<4, S>(row, sc);
        const uint32_t sx = h ? sc[2] : sc[0], sy = h ? sc[3] : sc[1];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const uint32_t b = ((j < 4 ? sx : sy) >> (8 * (j & 3))) & 0xFFu;
            make_sb(4.0f * d * (float)(b & 0xFu), dmin * (float)(b >> 4), s2[j], b2[j]);
        }
        lds_shifted<8, S>(row + 16 + 32 * h, qs);
    }
    __device__ __forceinline__ void emit(uint32_t, int, int t, uint32_t (&out)[32]) const {
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {   // 32 elements each: bit pair j = 2t + jj of every code byte
            const int j = 2 * t + jj;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                uint32_t o[4];
                const int si = 2 * j + (u >> 1);
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const uint32_t code = (((qs[2 * u + i] >> (2 * j)) & 0x03030303u) << 5) | 0x80808080u;
                    expand4(code, s2[si], b2[si], o[2 * i], o[2 * i + 1]);
                }
                { out[4 * ((4 * jj + u))] = o[0]; out[4 * ((4 * jj + u)) + 1] = o[1]; out[4 * ((4 * jj + u)) + 2] = o[2]; out[4 * ((4 * jj + u)) + 3] = o[3]; }
            }
        }
    }
};

// ---- Q3_K (round 2): hmask[32] | qs[64] | scales[12] (6 bit) | d; code q3 = 2 low bits | hmask bit (4n + j) << 2, value (q3 - 4) d (sc - 32)
//      = ((1 + q3/8) - 1.5) * 8 d (sc - 32).  110-byte blocks: window of 128 bytes, shift (110 kb) % 16.
template <> struct dq<T_Q3_K> {
    static constexpr int BB = 110;
    static constexpr int PITCH = 128;
    static constexpr int STEP16 = 55;
    static constexpr bool SHIFTED = true;
    uint32_t s2[8], hm[8], qs[8];
    template <int S> __device__ __forceinline__ void prep(uint32_t row, int h) {
        const float d = half_bits_to_float(lds16(row + 108 + S));
        uint32_t sc[3]; lds_shifted<3, S>(row + 96, sc);
#pragma unroll
        for (int j = 0; j < 8; ++j) {      // scale 8h + j: low nibble from scales[(8h + j) % 8] (high nibble for the second half), high pair from scales[8 + j % 4]
            const uint32_t lo = (sc[j >> 2] >> (8 * (j & 3) + 4 * h)) & 0xFu;
            const uint32_t hi = (sc[2] >> (8 * (j & 3) + 2 * (2 * h + (j >> 2)))) & 3u;
            s2[j] = pack_bf16_dup(8.0f * d * (float)((int)(lo | (hi << 4)) - 32));
        }
        lds_shifted<8, S>(row, hm);
        lds_shifted<8, S>(row + 32 + 32 * h, qs);
    }
    __device__ __forceinline__ void emit(uint32_t, int h, int t, uint32_t (&out)[32]) const {
#pragma unroll
        for (int jj = 0; jj < 2; ++jj) {
            const int j = 2 * t + jj;
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                uint32_t o[4];
                const int si = 2 * j + (u >> 1);
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const uint32_t code = (((qs[2 * u + i] >> (2 * j)) & 0x03030303u) << 4) | (((hm[2 * u + i] >> (4 * h + j)) & 0x01010101u) << 6) | 0x80808080u;
                    expand4_sym(code, s2[si], o[2 * i], o[2 * i + 1]);
                }
                { out[4 * ((4 * jj + u))] = o[0]; out[4 * ((4 * jj + u)) + 1] = o[1]; out[4 * ((4 * jj + u)) + 2] = o[2]; out[4 * ((4 * jj + u)) + 3] = o[3]; }
            }
        }
    }
};

// int8 x 4 (a word of table values) times an f16 scale -> two packed bf16 pairs.  As for Q8_0: 1024 + (v ^ 0x80) is exact in f16,
// minus 1152 gives v, times the scale, rounded to bf16.
__device__ __forceinline__ void i8x4_scaled(uint32_t v, __half2 d2, uint32_t & o01, uint32_t & o23) {
    const __half2 off = __float2half2_rn(1152.0f);
    const uint32_t x = v ^ 0x80808080u;
    const uint32_t a = __byte_perm(x, 0x64646464u, PRMT_01), c = __byte_perm(x, 0x64646464u, PRMT_23);
    const float2 fa = __half22float2(__hmul2(__hsub2(*reinterpret_cast<const __half2 *>(&a), off), d2));
    const float2 fc = __half22float2(__hmul2(__hsub2(*reinterpret_cast<const __half2 *>(&c), off), d2));
    o01 = pack_bf16(fa.x, fa.y); o23 = pack_bf16(fc.x, fc.y);
}

// ---- IQ4_NL (round 2): 8 blocks of {d (f16), qs[16]} per 256 K; table[nibble] * d (table: common.cuh iq4_lookup8)
template <> struct dq<T_IQ4_NL> {
    static constexpr int PITCH = 144;
    static constexpr int STEP16 = 72;
    static constexpr bool SHIFTED = false;
    template <int S> __device__ __forceinline__ void prep(uint32_t, int) {}
    __device__ __forceinline__ void emit(uint32_t row, int h, int t, uint32_t (&out)[32]) const {
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const uint32_t blk = row + 18 * (2 * (2 * h + t) + b);
            const __half2 d2 = __half2half2(__ushort_as_half((unsigned short)lds16(blk)));
            int w[4]; lds_words_a2<4>(blk + 2, w);
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                uint32_t o[4], p[4];
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    int vlo, vhi; iq4_lookup8((uint32_t)w[2 * u + i], vlo, vhi);
                    i8x4_scaled((uint32_t)vlo, d2, o[2 * i], o[2 * i + 1]);
                    i8x4_scaled((uint32_t)vhi, d2, p[2 * i], p[2 * i + 1]);
                }
                { out[4 * ((4 * b + u))] = o[0]; out[4 * ((4 * b + u)) + 1] = o[1]; out[4 * ((4 * b + u)) + 2] = o[2]; out[4 * ((4 * b + u)) + 3] = o[3]; }
                { out[4 * ((4 * b + u + 2))] = p[0]; out[4 * ((4 * b + u + 2)) + 1] = p[1]; out[4 * ((4 * b + u + 2)) + 2] = p[2]; out[4 * ((4 * b + u + 2)) + 3] = p[3]; }
            }
        }
    }
};

// ---- IQ4_XS (round 2): d | scales_h | scales_l[4] | qs[128]; block ib: table[nibble] * d * (ls - 32), ls = 6 bits.  136-byte blocks:
//      window of 144 bytes, shift (136 kb) % 16 = 0 / 8; the codes of a block sit 8 bytes further (shift 8 / 0 of the next 16 bytes).
template <> struct dq<T_IQ4_XS> {
    static constexpr int BB = 136;
    static constexpr int PITCH = 144;
    static constexpr int STEP16 = 68;
    static constexpr bool SHIFTED = true;
    uint32_t dl[4], qs[16];                // this half's four block scales (f16 pairs) and 64 code bytes
    template <int S> __device__ __forceinline__ void prep(uint32_t row, int h) {
        uint32_t hd[2]; lds_shifted<2, S>(row, hd);
        const float d = half_bits_to_float(hd[0] & 0xFFFFu);
        const uint32_t sh = hd[0] >> 16, sl = hd[1];
#pragma unroll
        for (int bb = 0; bb < 4; ++bb) {
            const int ib = 4 * h + bb;
            const int ls = (int)(((sl >> (4 * ib)) & 0xFu) | (((sh >> (2 * ib)) & 3u) << 4)) - 32;
            const __half2 d2 = __float2half2_rn(d * (float)ls);
            dl[bb] = *reinterpret_cast<const uint32_t *>(&d2);
            uint32_t w[4];
            lds_shifted<4, (S + 8) % 16>(row + 16 * ib + ((S + 8) & ~15), w);
            qs[4 * bb] = w[0]; qs[4 * bb + 1] = w[1]; qs[4 * bb + 2] = w[2]; qs[4 * bb + 3] = w[3];
        }
    }
    __device__ __forceinline__ void emit(uint32_t, int, int t, uint32_t (&out)[32]) const {
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const __half2 d2 = *reinterpret_cast<const __half2 *>(&dl[2 * t + b]);
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                uint32_t o[4], p[4];
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    int vlo, vhi; iq4_lookup8(qs[4 * (2 * t + b) + 2 * u + i], vlo, vhi);
                    i8x4_scaled((uint32_t)vlo, d2, o[2 * i], o[2 * i + 1]);
                    i8x4_scaled((uint32_t)vhi, d2, p[2 * i], p[2 * i + 1]);
                }
                { out[4 * ((4 * b + u))] = o[0]; out[4 * ((4 * b + u)) + 1] = o[1]; out[4 * ((4 * b + u)) + 2] = o[2]; out[4 * ((4 * b + u)) + 3] = o[3]; }
                { out[4 * ((4 * b + u + 2))] = p[0]; out[4 * ((4 * b + u + 2)) + 1] = p[1]; out[4 * ((4 * b + u + 2)) + 2] = p[2]; out[4 * ((4 * b + u + 2)) + 3] = p[3]; }
            }
        }
    }
};

// ---- Q4_0: 8 blocks of {d (f16), qs[16]} per 256 K; element j of a block: nibble (j/16) of qs[j%16]; (q - 8) * d
template <> struct dq<T_Q4_0> {
    static constexpr int PITCH = 144;
    static constexpr int STEP16 = 72;
    static constexpr bool SHIFTED = false;
    template <int S> __device__ __forceinline__ void prep(uint32_t, int) {}
    __device__ __forceinline__ void emit(uint32_t row, int h, int t, uint32_t (&out)[32]) const {
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const uint32_t blk = row + 18 * (2 * (2 * h + t) + b);
            const uint32_t s2 = pack_bf16_dup(16.0f * half_bits_to_float(lds16(blk)));      // (q - 8) d = ((1 + q/16) - 1.5) * 16d
            int w[4]; lds_words_a2<4>(blk + 2, w);
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                uint32_t o[4], p[4];
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const uint32_t x = (uint32_t)w[2 * u + i];
                    expand4_sym(((x << 3) & 0x78787878u) | 0x80808080u, s2, o[2 * i], o[2 * i + 1]);
                    expand4_sym(((x >> 1) & 0x78787878u) | 0x80808080u, s2, p[2 * i], p[2 * i + 1]);
                }
                { out[4 * ((4 * b + u))] = o[0]; out[4 * ((4 * b + u)) + 1] = o[1]; out[4 * ((4 * b + u)) + 2] = o[2]; out[4 * ((4 * b + u)) + 3] = o[3]; }
                { out[4 * ((4 * b + u + 2))] = p[0]; out[4 * ((4 * b + u + 2)) + 1] = p[1]; out[4 * ((4 * b + u + 2)) + 2] = p[2]; out[4 * ((4 * b + u + 2)) + 3] = p[3]; }
            }
        }
    }
};

// ---- Q4_1: 8 blocks of {d, m (f16), qs[16]} per 256 K; nibble * d + m = ((1 + q/16) - 1.5) * 16d + (8d + m)   (round 2)
template <> struct dq<T_Q4_1> {
    static constexpr int PITCH = 160;
    static constexpr int STEP16 = 80;
    static constexpr bool SHIFTED = false;
    template <int S> __device__ __forceinline__ void prep(uint32_t, int) {}
    __device__ __forceinline__ void emit(uint32_t row, int h, int t, uint32_t (&out)[32]) const {
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const uint32_t blk = row + 20 * (2 * (2 * h + t) + b);                          // 4-byte aligned
            const uint32_t dm = lds32(blk);
            uint32_t s2, c2;
            make_sb(16.0f * half_bits_to_float(dm & 0xFFFFu), -half_bits_to_float(dm >> 16), s2, c2);
            uint32_t w[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) w[j] = lds32(blk + 4 + 4 * j);
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                uint32_t o[4], p[4];
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const uint32_t x = w[2 * u + i];
                    expand4(((x << 3) & 0x78787878u) | 0x80808080u, s2, c2, o[2 * i], o[2 * i + 1]);
                    expand4(((x >> 1) & 0x78787878u) | 0x80808080u, s2, c2, p[2 * i], p[2 * i + 1]);
                }
                { out[4 * ((4 * b + u))] = o[0]; out[4 * ((4 * b + u)) + 1] = o[1]; out[4 * ((4 * b + u)) + 2] = o[2]; out[4 * ((4 * b + u)) + 3] = o[3]; }
                { out[4 * ((4 * b + u + 2))] = p[0]; out[4 * ((4 * b + u + 2)) + 1] = p[1]; out[4 * ((4 * b + u + 2)) + 2] = p[2]; out[4 * ((4 * b + u + 2)) + 3] = p[3]; }
            }
        }
    }
};

// four bits of x (bits 0..3) to bit 6 of four bytes: x * 0x00204081 puts bit k at bit 8k (the four shifted copies do not overlap)
__device__ __forceinline__ uint32_t spread4_bit6(uint32_t x) { return (((x & 0xFu) * 0x00204081u) & 0x01010101u) << 6; }

// ---- Q5_0 / Q5_1: 8 blocks of {d [, m], qh[4], qs[16]}; 5-bit code q = nibble | bit j of qh << 4 (element j; the high nibbles are
//      elements 16..31).  Q5_0: (q - 16) d = ((1 + q/32) - 1.5) * 32d;  Q5_1: q d + m = ((1 + q/32) - 1.5) * 32d + (16d + m).
//      Code byte 0x80 | q << 2: nibble at bits 2..5, fifth bit at bit 6.
template <int TYPE> struct dq5 {
    static constexpr int BB = TYPE == T_Q5_0 ? 22 : 24;
    static constexpr int PITCH = 8 * BB;
    static constexpr int STEP16 = 4 * BB;
    static constexpr bool SHIFTED = false;
    template <int S> __device__ __forceinline__ void prep(uint32_t, int) {}
    __device__ __forceinline__ void emit(uint32_t row, int h, int t, uint32_t (&out)[32]) const {
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const uint32_t blk = row + BB * (2 * (2 * h + t) + b);                          // Q5_0: 2-byte aligned, Q5_1: 8-byte aligned
            uint32_t s2, c2 = 0, qh, w[4];
            if constexpr (TYPE == T_Q5_0) {
                s2 = pack_bf16_dup(32.0f * half_bits_to_float(lds16(blk)));
                int v[5]; lds_words_a2<5>(blk + 2, v);
                qh = (uint32_t)v[0];
#pragma unroll
                for (int j = 0; j < 4; ++j) w[j] = (uint32_t)v[1 + j];
            } else {
                const uint32_t dm = lds32(blk);
                make_sb(32.0f * half_bits_to_float(dm & 0xFFFFu), -half_bits_to_float(dm >> 16), s2, c2);
                qh = lds32(blk + 4);
#pragma unroll
                for (int j = 0; j < 4; ++j) w[j] = lds32(blk + 8 + 4 * j);
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                uint32_t o[4], p[4];
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const int e = 4 * (2 * u + i);                                          // first element of this word's low nibbles
                    const uint32_t x = w[2 * u + i];
                    const uint32_t lo = ((x << 2) & 0x3C3C3C3Cu) | 0x80808080u | spread4_bit6(qh >> e);
                    const uint32_t hi = ((x >> 2) & 0x3C3C3C3Cu) | 0x80808080u | spread4_bit6(qh >> (16 + e));
                    if constexpr (TYPE == T_Q5_0) { expand4_sym(lo, s2, o[2 * i], o[2 * i + 1]); expand4_sym(hi, s2, p[2 * i], p[2 * i + 1]); }
                    else { expand4(lo, s2, c2, o[2 * i], o[2 * i + 1]); expand4(hi, s2, c2, p[2 * i], p[2 * i + 1]); }
                }
                { out[4 * ((4 * b + u))] = o[0]; out[4 * ((4 * b + u)) + 1] = o[1]; out[4 * ((4 * b + u)) + 2] = o[2]; out[4 * ((4 * b + u)) + 3] = o[3]; }
                { out[4 * ((4 * b + u + 2))] = p[0]; out[4 * ((4 * b + u + 2)) + 1] = p[1]; out[4 * ((4 * b + u + 2)) + 2] = p[2]; out[4 * ((4 * b + u + 2)) + 3] = p[3]; }
            }
        }
    }
};
template <> struct dq<T_Q5_0> : dq5<T_Q5_0> {};
template <> struct dq<T_Q5_1> : dq5<T_Q5_1> {};

// ---- Q8_0: 8 blocks of {d (f16), qs int8[32]} per 256 K; q * d.  Codes go through f16 (1024 + (q ^ 0x80) is
//      exact, minus 1152 gives q, times the f16 scale) and are then rounded to bf16.
template <> struct dq<T_Q8_0> {
    static constexpr int PITCH = 272;
    static constexpr int STEP16 = 136;
    static constexpr bool SHIFTED = false;
    template <int S> __device__ __forceinline__ void prep(uint32_t, int) {}
    __device__ __forceinline__ void emit(uint32_t row, int h, int t, uint32_t (&out)[32]) const {
#pragma unroll
        for (int b = 0; b < 2; ++b) {
            const uint32_t blk = row + 34 * (2 * (2 * h + t) + b);
            const __half dh = __ushort_as_half((unsigned short)lds16(blk));
            const __half2 d2 = __half2half2(dh), off = __float2half2_rn(1152.0f);
            int w[8]; lds_words_a2<8>(blk + 2, w);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                uint32_t o[4];
#pragma unroll
                for (int i = 0; i < 2; ++i) {
                    const uint32_t x = (uint32_t)w[2 * u + i] ^ 0x80808080u;
                    const uint32_t a = __byte_perm(x, 0x64646464u, PRMT_01), c = __byte_perm(x, 0x64646464u, PRMT_23);
                    const float2 fa = __half22float2(__hmul2(__hsub2(*reinterpret_cast<const __half2 *>(&a), off), d2));
                    const float2 fc = __half22float2(__hmul2(__hsub2(*reinterpret_cast<const __half2 *>(&c), off), d2));
                    o[2 * i] = pack_bf16(fa.x, fa.y); o[2 * i + 1] = pack_bf16(fc.x, fc.y);
                }
                { out[4 * ((4 * b + u))] = o[0]; out[4 * ((4 * b + u)) + 1] = o[1]; out[4 * ((4 * b + u)) + 2] = o[2]; out[4 * ((4 * b + u)) + 3] = o[3]; }
            }
        }
    }
};

// ------------------------------------------------------------------------------------------------
// The kernel.  Step i = 4*kb + j covers the 64-K sub-chunk c = {0,2,1,3}[j] of block kb; dequantiser group g (four warps,
// one per TMEM lane quarter) owns sub-chunk c = g of every block, i.e. step 4*kb + 2*(g & 1) + (g >> 1).
// Tensor memory: columns [0, tn) hold the f32 accumulator, columns [a_col0, a_col0 + 32*na) the ring of
// `na` A slots (128 lanes x 32 columns = 128 rows x 64 bf16 each).
// ------------------------------------------------------------------------------------------------
struct mmq_params {
    float * dst; int64_t ds;     // dst[m*ds + n]
    int N, M, K;
    int tn;                      // tokens per tile: 32, 64, 128 or 256
    int nb;                      // depth of the activation-tile ring in shared memory (<= NB_MAX)
    int na;                      // depth of the A ring in tensor memory: 4 or 8 (even)
    int nr;                      // depth of the raw weight-block ring in shared memory: 2 or 4
    int tmem_cols;               // 256 or 512
    long long * trace;           // debug: clock64 stamps of CTA (0,0), [event][index] (b200_mmq_debug_trace)
    // MUL_MAT_ID (expert = blockIdx.z): activation rows sorted by expert; rows [off[e], off[e] + cnt[e]) belong to expert e,
    // row m of the sorted list goes to dst + row_off[m]; the weight map stacks the experts' matrices (N rows each)
    const int * moe_cnt; const int * moe_off; const int64_t * moe_row;
    // grouped launch (matrices of one type and K that read the same activations, e.g. q/k/v): group g owns the row tiles
    // [gy[g], gy[g+1]) of grid.y, its own tensor map (wmap / wmap1 / wmap2), destination and row count
    int ngroup; int gy[4]; float * gdst[3]; int64_t gds[3]; int gN[3];
    // split along K for launches of one token tile (M <= tn: batched decode, short prompts), whose grid would otherwise be N / 128
    // CTAs (32 for a 4096-row matrix on 148 SMs): CTA z multiplies K blocks [z * nkb, (z + 1) * nkb) into part[z][M][part_ds]
    // (group g at column gcol[g]); ksplit_reduce_kernel adds the slices in fixed order.  ksplit <= 1: off.
    // In a grouped launch every matrix keeps the slice count its own launch would have (gks[g]; CTAs with z >= gks[g] leave at
    // once): grouped and single launches add in the same order, so fused and unfused graphs stay bit-identical.
    int ksplit; float * part; int part_ds; int gcol[3]; int gks[3];
};
constexpr int TRACE_N = 512;
#ifdef B200_MMQ_TRACE
__device__ __forceinline__ void trace(const mmq_params & p, int ev, int idx) {
    if (p.trace && blockIdx.x == 0 && blockIdx.y == 0 && idx < TRACE_N) p.trace[ev * TRACE_N + idx] = clock64();
}
#else
__device__ __forceinline__ void trace(const mmq_params &, int, int) {}
#endif

__device__ __forceinline__ int by_group(const mmq_params & p, int by) { return by >= p.gy[2] ? 2 : by >= p.gy[1] ? 1 : 0; }

constexpr int MMQ_THREADS = 640;            // TMA x2, MMA issue, TMEM allocator warps + 16 dequantiser warps
constexpr int NA_MAX = 8;
constexpr int NB_MAX = 6;

template <int TYPE>
__global__ void __launch_bounds__(MMQ_THREADS, 1)
mmq_kernel(const __grid_constant__ CUtensorMap wmap0, const __grid_constant__ CUtensorMap xmap, const mmq_params p,
           const __grid_constant__ CUtensorMap wmap1, const __grid_constant__ CUtensorMap wmap2) {
    using DQ = dq<TYPE>;
    constexpr int RAW = 128 * DQ::PITCH;
    extern __shared__ uint8_t smem_dyn[];
    const uint32_t base = (smem_u32(smem_dyn) + 1023u) & ~1023u;
    const int tn = p.tn, nb = p.nb, na = p.na;
    const uint32_t b_slot = (uint32_t)tn * 128u;
    const int nr = p.nr, lg_nr = nr == 4 ? 2 : 1;
    const uint32_t sB = base, sR = sB + (uint32_t)nb * b_slot, sBar = sR + (uint32_t)nr * RAW;
    const uint32_t raw_full = sBar, raw_empty = sBar + 32, a_full = sBar + 64, a_empty = a_full + 8 * NA_MAX,
                   b_full = a_empty + 8 * NA_MAX, b_empty = b_full + 8 * NB_MAX, tmem_full = b_empty + 8 * NB_MAX, tmem_slot = tmem_full + 8;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const CUtensorMap * wm = &wmap0;
    float * dst_g = p.dst; int64_t ds_g = p.ds; int N_g = p.N; int by = blockIdx.y; int gcol = 0;
    if (p.ngroup > 1) {
        const int g = by >= p.gy[2] ? 2 : by >= p.gy[1] ? 1 : 0;
        by -= p.gy[g];
        wm = g == 0 ? &wmap0 : g == 1 ? &wmap1 : &wmap2;
        dst_g = p.gdst[g]; ds_g = p.gds[g]; N_g = p.gN[g]; gcol = p.gcol[g];
    }
    const int ks = p.ksplit > 1 ? (p.ngroup > 1 ? p.gks[by_group(p, blockIdx.y)] : p.ksplit) : 1;
    if (p.ksplit > 1 && !p.moe_cnt && (int)blockIdx.z >= ks) return;      // (before any barrier: the whole CTA leaves)
    if (ks > 1) { dst_g = p.part + (int64_t)blockIdx.z * p.M * p.part_ds + gcol; ds_g = p.part_ds; }
    int m0 = blockIdx.x * tn, m_end = p.M, w_row0 = by * 128;
    const int n0 = by * 128;
    if (p.moe_cnt) {                             // (launched without PDL: the routing tables are complete)
        const int cnt = p.moe_cnt[blockIdx.z];
        if (m0 >= cnt) return;                   // this expert has fewer tokens than the worst case the grid was sized for
        m0 += p.moe_off[blockIdx.z];
        m_end = p.moe_off[blockIdx.z] + cnt;
        w_row0 += blockIdx.z * p.N;
    }
    const int nkb = (p.K >> 8) / ks, nsteps = nkb * 4;      // this CTA's K blocks: [kb0, kb0 + nkb)
    const int kb0 = ks > 1 ? (int)blockIdx.z * nkb : 0;
    const uint32_t a_col0 = (uint32_t)p.tmem_cols - 32u * (uint32_t)na;

    if (threadIdx.x == 0) {
        for (int i = 0; i < 4; ++i) { mbar_init(raw_full + 8 * i, 1); mbar_init(raw_empty + 8 * i, 16); }
        for (int i = 0; i < NA_MAX; ++i) { mbar_init(a_full + 8 * i, 4); mbar_init(a_empty + 8 * i, 1); }
        for (int i = 0; i < NB_MAX; ++i) { mbar_init(b_full + 8 * i, 1); mbar_init(b_empty + 8 * i, 1); }
        mbar_init(tmem_full, 1);
        mbar_fence_init();
        tma_prefetch_desc(wm); tma_prefetch_desc(&xmap);
    }
    if (warp == 2) { tmem_alloc(tmem_slot, (uint32_t)p.tmem_cols); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = lds32(tmem_slot);
    pdl_trigger();
    pdl_wait();                                  // the bf16 activations come from the preceding kernel

    if (warp == 0) {
        if (lane == 0) {                         // raw weight blocks: one 128-row box per 256-K block
            for (int kb = 0; kb < nkb; ++kb) {
                const int s = kb & (nr - 1);
                mbar_wait_b(raw_empty + 8 * s, ((kb >> lg_nr) & 1) ^ 1);
                trace(p, 0, kb);
                mbar_expect_tx(raw_full + 8 * s, RAW);
                tma_load_2d(sR + s * RAW, wm, ((kb0 + kb) * DQ::STEP16) & ~7, w_row0, raw_full + 8 * s);     // 16-byte aligned start
            }
        }
    } else if (warp == 3) {
        if (lane == 0) {                         // bf16 activation tiles [tn][64], 128-byte swizzle
            int s = 0; uint32_t ph = 1;
            for (int i = 0; i < nsteps; ++i) {
                const int j = i & 3, c = ((j & 1) << 1) | (j >> 1);
                mbar_wait(b_empty + 8 * s, ph);
                trace(p, 1, i);
                mbar_expect_tx(b_full + 8 * s, b_slot);
                tma_load_2d(sB + s * b_slot, &xmap, (kb0 + (i >> 2)) * 256 + c * 64, m0, b_full + 8 * s);
                if (++s == nb) { s = 0; ph ^= 1; }
            }
        }
    } else if (warp == 1) {
        if (elect_one()) {                       // tensor-core issue: one thread, so every instruction between two MMAs counts
            const uint32_t idesc = umma_idesc(1, tn, 0);
            const uint32_t a0 = tmem_base + a_col0;
            const uint64_t bd0 = umma_desc_sw128(sB);
            const uint32_t bd_step = b_slot >> 4;
            int sa = 0, sb = 0; uint32_t pa = 0, pb = 0, acc = 0;
            for (int i = 0; i < nsteps; ++i) {
                trace(p, 2, i);
                mbar_wait(a_full + 8 * sa, pa);
                mbar_wait(b_full + 8 * sb, pb);
                trace(p, 4, i);
                tc_fence_after();
                const uint32_t at = a0 + 32u * (uint32_t)sa;
                const uint64_t bd = bd0 + (uint64_t)(bd_step * (uint32_t)sb);
                tc_mma_f16_ts(tmem_base, at, bd, idesc, acc);      // K = 16 per instruction: 8 TMEM columns of A, 32 bytes along B's swizzled row
                tc_mma_f16_ts(tmem_base, at + 8, bd + 2, idesc, 1);
                tc_mma_f16_ts(tmem_base, at + 16, bd + 4, idesc, 1);
                tc_mma_f16_ts(tmem_base, at + 24, bd + 6, idesc, 1);
                acc = 1;
                tc_commit(a_empty + 8 * sa);
                tc_commit(b_empty + 8 * sb);
                trace(p, 5, i);
                if (++sa == na) { sa = 0; pa ^= 1; }
                if (++sb == nb) { sb = 0; pb ^= 1; }
            }
            tc_commit(tmem_full);
        }
    } else if (warp >= 4) {
        // sixteen dequantiser warps: warp quarter (dw & 3) = 32 weight rows = its TMEM lane quarter, group g = dw >> 2 = the 64-K
        // sub-chunk (h = g >> 1, t = g & 1) it expands in every block.  (Eight warps doing two sub-chunks each left the kernel
        // bound by their issue rate for tiles of <= 128 tokens: 1800 cycles per block against 1024 of tensor-pipe time.)
        const int dw = warp - 4, g = dw >> 2, h = g >> 1, r = (dw & 3) * 32 + lane;
        const uint32_t lane_base = tmem_base + ((uint32_t)((dw & 3) * 32) << 16);      // this warp's TMEM lane quarter
        const int na_shift = na == 8 ? 3 : 2;
        auto run = [&](auto tc) {
            constexpr int t = decltype(tc)::value;
            DQ q;
            auto block = [&](auto shift, int kb) {
                constexpr int S = decltype(shift)::value;
                const int s = kb & (nr - 1);
                const bool tr = lane == 0 && (dw & 3) == 0;
                if (tr && t == 0) trace(p, 8 + 8 * h, kb);
                mbar_wait_b(raw_full + 8 * s, (kb >> lg_nr) & 1);
                if (tr && t == 0) trace(p, 9 + 8 * h, kb);
                const uint32_t row = sR + s * RAW + r * DQ::PITCH;
                q.template prep<S>(row, h);
                const int i = 4 * kb + 2 * t + h, slot = i & (na - 1);
                uint32_t out[32];
                q.emit(row, h, t, out);
                if (tr) trace(p, 10 + 8 * h + 3 * t, kb);
                mbar_wait_b(a_empty + 8 * slot, ((i >> na_shift) & 1) ^ 1);
                if (tr) trace(p, 11 + 8 * h + 3 * t, kb);
                tc_fence_after();
                tmem_st32(lane_base + a_col0 + 32u * (uint32_t)slot, out);
                tmem_st_wait();
                tc_fence_before();
                if (tr) trace(p, 12 + 8 * h + 3 * t, kb);
                __syncwarp();
                if (lane == 0) { mbar_arrive(a_full + 8 * slot); mbar_arrive(raw_empty + 8 * s); }
            };
            if constexpr (DQ::SHIFTED) {         // block kb sits (BB * kb) % 16 bytes into its raw row (Q6_K: 2 * (kb % 8)): unroll the period of 8
                constexpr int BBs = DQ::BB;
                for (int kb = 0; kb < nkb; kb += 8) {
                    block(std::integral_constant<int, 0>{}, kb);                    block(std::integral_constant<int, (BBs * 1) % 16>{}, kb + 1);
                    block(std::integral_constant<int, (BBs * 2) % 16>{}, kb + 2);  block(std::integral_constant<int, (BBs * 3) % 16>{}, kb + 3);
                    block(std::integral_constant<int, (BBs * 4) % 16>{}, kb + 4);  block(std::integral_constant<int, (BBs * 5) % 16>{}, kb + 5);
                    block(std::integral_constant<int, (BBs * 6) % 16>{}, kb + 6);  block(std::integral_constant<int, (BBs * 7) % 16>{}, kb + 7);
                }
            } else {
                for (int kb = 0; kb < nkb; ++kb) block(std::integral_constant<int, 0>{}, kb);
            }
        };
        if (g & 1) run(std::integral_constant<int, 1>{}); else run(std::integral_constant<int, 0>{});
        // epilogue: TMEM lane = weight row (this warp's quarter), column = token; group g takes every fourth 32 columns
        mbar_wait_b(tmem_full, 0);
        tc_fence_after();
        if (lane == 0 && dw == 0) trace(p, 6, 0);
        // (addresses are stepped, not recomputed: with the row-table test inside the store loop every store paid a chain of
        // constant loads and 64-bit multiplies, and the epilogue was a fifth of the kernel at K = 4096)
        const int n = n0 + r;
        float * const dbase = dst_g; const int64_t ds = ds_g; const int64_t * const mrow = p.moe_row;
        for (int c0 = g * 32; c0 < tn; c0 += 128) {
            uint32_t v[32];
            tmem_ld32(lane_base + (uint32_t)c0, v);
            tmem_ld_wait();
            const int mb = m0 + c0, cnt = m_end - mb;          // tokens of this column group that exist
            if (n < N_g && cnt > 0) {
                if (mrow) {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (j < cnt) dbase[mrow[mb + j] + n] = __uint_as_float(v[j]);
                } else {
                    float * d = dbase + (int64_t)mb * ds + n;
                    if (cnt >= 32) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) { *d = __uint_as_float(v[j]); d += ds; }
                    } else {
#pragma unroll
                        for (int j = 0; j < 32; ++j) { if (j < cnt) *d = __uint_as_float(v[j]); d += ds; }
                    }
                }
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x == 0) trace(p, 7, 0);
    if (warp == 2) { tc_fence_after(); tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols); }
}

// ---- activations f32 -> bf16 [M][K] (contiguous) ----
__global__ void __launch_bounds__(256) x_to_bf16_kernel(const float * __restrict__ X, int64_t xs, uint32_t * __restrict__ out, int K, int M, int vec) {
    pdl_trigger(); pdl_wait();
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;      // one group of 8 elements
    const int kg = K >> 3;
    if (i >= (int64_t)M * kg) return;
    const int m = (int)(i / kg), k = (int)(i % kg) * 8;
    const float * src = X + (int64_t)m * xs + k;
    float f[8];
    if (vec) {
        const float4 a = *reinterpret_cast<const float4 *>(src), b = *reinterpret_cast<const float4 *>(src + 4);
        f[0] = a.x; f[1] = a.y; f[2] = a.z; f[3] = a.w; f[4] = b.x; f[5] = b.y; f[6] = b.z; f[7] = b.w;
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) f[j] = src[j];
    }
    uint4 o; o.x = pack_bf16(f[0], f[1]); o.y = pack_bf16(f[2], f[3]); o.z = pack_bf16(f[4], f[5]); o.w = pack_bf16(f[6], f[7]);
    *reinterpret_cast<uint4 *>(out + ((int64_t)m * K + k) / 2) = o;
}

// ---- MUL_MAT_ID routing (reference ggml-cuda/ggml-cuda.cu:2010-2140 sorts the rows on the host after a
// device->host copy of ids and a stream synchronisation; here the tables never leave the device) ----
// One CTA: counts per expert, exclusive prefix, then the (token, slot) pairs in expert order.
__global__ void __launch_bounds__(1024) moe_sort_kernel(const int32_t * __restrict__ ids, int64_t ids_s1, int n_used, int T, int n_expert,
                                                        int * __restrict__ cnt, int * __restrict__ off, int2 * __restrict__ pairs) {
    extern __shared__ int sm_moe[];              // [n_expert] counters, [n_expert] cursors
    int * c = sm_moe, * cur = sm_moe + n_expert;
    pdl_trigger(); pdl_wait();
    for (int e = threadIdx.x; e < n_expert; e += blockDim.x) c[e] = 0;
    __syncthreads();
    const int total = n_used * T;
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
        const int e = ids[(int64_t)(i / n_used) * ids_s1 + i % n_used];
        if (e >= 0 && e < n_expert) atomicAdd(&c[e], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int run = 0;
        for (int e = 0; e < n_expert; ++e) { cur[e] = run; off[e] = run; cnt[e] = c[e]; run += c[e]; }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < total; i += blockDim.x) {
        const int t = i / n_used, u = i % n_used;
        const int e = ids[(int64_t)t * ids_s1 + u];
        if (e >= 0 && e < n_expert) pairs[atomicAdd(&cur[e], 1)] = make_int2(t, u);
    }
}
// One CTA per sorted row: its activations f32 -> bf16, and where its result goes.
__global__ void __launch_bounds__(256) moe_gather_kernel(const float * __restrict__ X, int64_t x_sy, int64_t x_sz, int x_ne1, const int2 * __restrict__ pairs,
                                                         const int * __restrict__ cnt, const int * __restrict__ off, int n_expert,
                                                         uint32_t * __restrict__ out, int64_t * __restrict__ row_off, int64_t d_sy, int64_t d_sz, int K) {
    const int slot = blockIdx.x;
    if (slot >= off[n_expert - 1] + cnt[n_expert - 1]) return;        // ids outside [0, n_expert) leave the tail unused
    const int2 tu = pairs[slot];
    const float * src = X + (int64_t)tu.x * x_sz + (int64_t)(tu.y % x_ne1) * x_sy;
    for (int k = threadIdx.x * 2; k < K; k += 512) out[((int64_t)slot * K + k) >> 1] = pack_bf16(src[k], src[k + 1]);
    if (threadIdx.x == 0) row_off[slot] = (int64_t)tu.x * d_sz + (int64_t)tu.y * d_sy;
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
encode_tiled_fn encode_tiled() {
    static encode_tiled_fn fn = [] {
        void * p = nullptr; cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) p = nullptr;
        return (encode_tiled_fn)p;
    }();
    return fn;
}

static int env_int(const char * name, int def) { const char * e = getenv(name); return e ? atoi(e) : def; }

template <int TYPE> static int pitch_of() { return dq<TYPE>::PITCH; }
// Q2_K / Q3_K / IQ4_NL / IQ4_XS on the tensor-core path (round 2; B200_MMQ_MORE=0: those types run prefill as mat-vec passes)
static bool mmq_more() { static const bool v = [] { const char * e = getenv("B200_MMQ_MORE"); return !e || atoi(e) != 0; }(); return v; }
// formats whose blocks are not multiples of 16 bytes: shifted windows, whole periods of 8 blocks per row
static bool mmq_shifted(int type) { return type == T_Q6_K || type == T_Q2_K || type == T_Q3_K || type == T_IQ4_XS; }
static int raw_pitch(int type) {
    switch (type) { case T_Q4_0: return pitch_of<T_Q4_0>(); case T_Q8_0: return pitch_of<T_Q8_0>(); case T_Q4_K: return pitch_of<T_Q4_K>();
                    case T_Q4_1: return pitch_of<T_Q4_1>(); case T_Q5_0: return pitch_of<T_Q5_0>(); case T_Q5_1: return pitch_of<T_Q5_1>();
                    case T_Q5_K: return pitch_of<T_Q5_K>(); case T_Q6_K: return pitch_of<T_Q6_K>();
                    case T_Q2_K: return mmq_more() ? pitch_of<T_Q2_K>() : 0; case T_Q3_K: return mmq_more() ? pitch_of<T_Q3_K>() : 0;
                    case T_IQ4_NL: return mmq_more() ? pitch_of<T_IQ4_NL>() : 0; case T_IQ4_XS: return mmq_more() ? pitch_of<T_IQ4_XS>() : 0; }
    return 0;
}
static size_t mmq_smem(int type, int tn, int nb, int nr) { return 1024 + (size_t)nb * tn * 128 + (size_t)nr * 128 * (size_t)raw_pitch(type) + 512; }
// ring depths: a deeper raw-block ring (more weight bytes in flight per SM) is worth more than the last activation slots
static void choose_rings(int type, mmq_params & p) {
    static const int want_nr = env_int("B200_MMQ_NR", 4) >= 4 ? 4 : 2, min_nb = std::max(2, std::min(NB_MAX, env_int("B200_MMQ_MIN_NB", 3)));
    p.nr = want_nr; p.nb = NB_MAX;
    while (p.nb > min_nb && mmq_smem(type, p.tn, p.nb, p.nr) > 232448) --p.nb;
    if (mmq_smem(type, p.tn, p.nb, p.nr) > 232448) { p.nr = 2; p.nb = NB_MAX; }
    while (p.nb > 2 && mmq_smem(type, p.tn, p.nb, p.nr) > 232448) --p.nb;
}

// tokens per tile: fill the 148 SMs first, then prefer the wider tile (weights are dequantised once per tile)
static int choose_tn(int N, int M) {
    const int forced = env_int("B200_MMQ_TN", 0);
    if (forced == 32 || forced == 64 || forced == 128 || forced == 256) return forced;
    if (M <= 32) return 32;
    if (M <= 64) return 64;
    if (M <= 128) return 128;
    const int rows = (N + 127) / 128;
    double best = 1e30; int best_tn = 128;
    for (int tn : {256, 128}) {
        const int tiles = rows * ((M + tn - 1) / tn);
        const double cost = (double)((tiles + 147) / 148) * (tn == 256 ? 2.0 : 1.25);      // per-tile time: tensor pipe, or dequantiser for narrow tiles
        if (cost < best) { best = cost; best_tn = tn; }
    }
    return best_tn;
}

template <int TYPE>
static int mmq_launch(const CUtensorMap & wmap, const CUtensorMap & xmap, const mmq_params & p, size_t smem, cudaStream_t st, int n_expert = 0,
                      const CUtensorMap * wmap1 = nullptr, const CUtensorMap * wmap2 = nullptr) {
    static size_t configured[16] = {0};  // per device
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    if (smem > configured[dev & 15]) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(mmq_kernel<TYPE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev & 15] = smem;
    }
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((p.M + p.tn - 1) / p.tn, p.ngroup > 1 ? p.gy[p.ngroup] : (p.N + 127) / 128, n_expert > 0 ? n_expert : std::max(1, p.ksplit)); lc.blockDim = dim3(MMQ_THREADS); lc.dynamicSmemBytes = smem; lc.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = pdl_default() && n_expert == 0 ? 1 : 0;       // the expert tables are read before griddepcontrol.wait
    B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, mmq_kernel<TYPE>, wmap, xmap, p, wmap1 ? *wmap1 : wmap, wmap2 ? *wmap2 : wmap));
    ++g_launches;
    return 0;
}

// ---- split along K for one-token-tile launches ----
__global__ void __launch_bounds__(256) ksplit_reduce_kernel(const float * __restrict__ part, int ks0, int ks1, int ks2, int M, int part_ds, int ng,
                                                            int c0, int c1, int c2, int n0, int n1, int n2,
                                                            float * d0, float * d1, float * d2, int64_t s0, int64_t s1, int64_t s2) {
    pdl_trigger(); pdl_wait();
    const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (i >= (int64_t)M * part_ds) return;
    const int m = (int)(i / part_ds), c = (int)(i % part_ds);
    const int g = ng > 2 && c >= c2 ? 2 : ng > 1 && c >= c1 ? 1 : 0;
    const int n = c - (g == 0 ? c0 : g == 1 ? c1 : c2);
    if (n >= (g == 0 ? n0 : g == 1 ? n1 : n2)) return;
    const int ks = g == 0 ? ks0 : g == 1 ? ks1 : ks2;
    if (ks <= 1) return;                          // this matrix was not split: its CTAs wrote dst themselves
    float v = 0.0f;
    for (int z = 0; z < ks; ++z) v += part[((int64_t)z * M + m) * part_ds + c];      // fixed order: deterministic
    float * d = g == 0 ? d0 : g == 1 ? d1 : d2;
    d[(int64_t)m * (g == 0 ? s0 : g == 1 ? s1 : s2) + n] = v;
}
// One buffer per device, allocated once (a replayed CUDA graph keeps its address) at the first use outside stream capture.
constexpr size_t KSPLIT_BYTES = (size_t)64 << 20;
static float * ksplit_buffer(cudaStream_t st) {
    static float * buf[16] = {nullptr};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return nullptr;
    if (!buf[dev]) {
        cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
        if (cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) return nullptr;
        if (cudaMalloc((void **)&buf[dev], KSPLIT_BYTES) != cudaSuccess) { cudaGetLastError(); buf[dev] = nullptr; }
    }
    return buf[dev];
}
// how many slices: fill the SMs, keep at least 4 blocks per CTA, Q6_K slices in whole 8-block alignment periods
static int plan_ksplit(int type, int K, int M, int tn, int ctas, int part_ds, cudaStream_t st, float ** part) {
    static const int on = env_int("B200_MMQ_KSPLIT", 1);
    *part = nullptr;
    if (!on || M > tn || M > 32 || ctas <= 0 || ctas > 74) return 1;      // (one 32-token tile: batched decode steps, short prompts)
    const int nkb = K >> 8, unit = mmq_shifted(type) ? 8 : 1;
    int ks = std::min(8, 148 / ctas);
    while (ks > 1 && (nkb % (ks * unit) != 0 || nkb / ks < 4)) --ks;
    if (ks < 2 || (size_t)ks * M * part_ds * 4 > KSPLIT_BYTES) return 1;
    *part = ksplit_buffer(st);
    return *part ? ks : 1;
}
static int ksplit_reduce(const mmq_params & p, cudaStream_t st) {
    const int ng = std::max(1, p.ngroup);
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)(((int64_t)p.M * p.part_ds + 255) / 256)); lc.blockDim = dim3(256); lc.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = pdl_default() ? 1 : 0;
    if (ng == 1) B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, ksplit_reduce_kernel, (const float *)p.part, p.ksplit, 0, 0, p.M, p.part_ds, 1, 0, 0, 0, p.N, 0, 0,
                                                    p.dst, (float *)nullptr, (float *)nullptr, p.ds, (int64_t)0, (int64_t)0));
    else B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, ksplit_reduce_kernel, (const float *)p.part, p.gks[0], p.gks[1], p.gks[2], p.M, p.part_ds, ng, p.gcol[0], p.gcol[1], p.gcol[2],
                                            p.gN[0], p.gN[1], p.gN[2], p.gdst[0], p.gdst[1], p.gdst[2], p.gds[0], p.gds[1], p.gds[2]));
    ++g_launches;
    return 0;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_mul_mat_q_supported(int type, const void * W, int64_t nb1, int K) {
    if (!raw_pitch(type) || K <= 0 || K % 256) return 0;
    if (mmq_shifted(type) && K % 2048) return 0;                  // the 210-byte (84, 110, 136) blocks realign with 16 bytes every 8 blocks at the latest
    if (((uintptr_t)W & 15) || (nb1 & 15)) return 0;              // TMA: 16-byte aligned base and row stride
    if (nb1 < (int64_t)(K / block_elems(type)) * block_bytes(type)) return 0;
    return encode_tiled() != nullptr;
}

static long long * g_mmq_trace = nullptr;
extern "C" int b200_mmq_debug_trace(void * dev_buf) {
#ifdef B200_MMQ_TRACE
    g_mmq_trace = (long long *)dev_buf; return 24 * TRACE_N;
#else
    (void)dev_buf; return 0;      // compiled without -DB200_MMQ_TRACE
#endif
}

extern "C" int64_t b200_mul_mat_q_scratch_bytes(int K, int M) { return (int64_t)K * M * 2 + 256; }

extern "C" int64_t b200_acts_bf16_bytes(int K, int M) { return (int64_t)K * M * 2; }

// activations f32 [M][xs] -> bf16 [M][K] contiguous (out 128-byte aligned)
extern "C" int b200_convert_acts_bf16(const float * X, int64_t xs, void * out, int K, int M, void * stream) {
    if (K <= 0 || M <= 0 || K % 8 || ((uintptr_t)out & 127)) return -1;
    const int vec = (((uintptr_t)X & 15) == 0 && (xs & 3) == 0) ? 1 : 0;
    const int64_t groups = (int64_t)M * (K / 8);
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((unsigned)((groups + 255) / 256)); lc.blockDim = dim3(256); lc.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = pdl_default() ? 1 : 0;
    B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, x_to_bf16_kernel, X, xs, (uint32_t *)out, K, M, vec));
    ++g_launches;
    return 0;
}

// dst = W x Xb^T with Xb the bf16 [M][K] activations made by b200_convert_acts_bf16
extern "C" int b200_mul_mat_q_bf16(int type, const void * W, int64_t nb1, const void * Xb, float * dst, int64_t ds,
                                   int N, int K, int M, void * stream) {
    if (!b200_mul_mat_q_supported(type, W, nb1, K) || N <= 0 || M <= 0 || ((uintptr_t)Xb & 127)) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    uint8_t * xb = (uint8_t *)Xb;

    mmq_params p = {};
    p.dst = dst; p.ds = ds; p.N = N; p.M = M; p.K = K;
    p.tn = choose_tn(N, M);
    p.trace = g_mmq_trace;
    choose_rings(type, p);
    // tensor memory: accumulator columns + A ring; narrow tiles take half of it so that two CTAs can share an SM
    p.na = p.tn <= 128 ? 4 : NA_MAX;
    p.tmem_cols = p.tn <= 128 ? 256 : 512;
    if (mmq_smem(type, p.tn, p.nb, p.nr) > 232448) return -3;
    const size_t smem = mmq_smem(type, p.tn, p.nb, p.nr);

    CUtensorMap wmap, xmap;
    {   // weights as a [N][nb1/2] array of 16-bit words; box = one 256-K block (padded to the raw pitch) x 128 rows
        const cuuint64_t dims[2] = {(cuuint64_t)((int64_t)(K / block_elems(type)) * block_bytes(type) / 2), (cuuint64_t)N};
        const cuuint64_t strides[1] = {(cuuint64_t)nb1};
        const cuuint32_t box[2] = {(cuuint32_t)(raw_pitch(type) / 2), 128};
        const cuuint32_t es[2] = {1, 1};
        if (encode_tiled()(&wmap, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, const_cast<void *>(W), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) return -4;
    }
    {   // activations bf16 [M][K]; box = 64 K x tn tokens, 128-byte swizzle
        const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)M};
        const cuuint64_t strides[1] = {(cuuint64_t)K * 2};
        const cuuint32_t box[2] = {64, (cuuint32_t)p.tn};
        const cuuint32_t es[2] = {1, 1};
        if (encode_tiled()(&xmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, xb, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) return -5;
    }
    p.part_ds = N;
    p.ksplit = plan_ksplit(type, K, M, p.tn, (N + 127) / 128, p.part_ds, st, &p.part);
    int rc = -1;
    switch (type) {
        case T_Q4_0: rc = mmq_launch<T_Q4_0>(wmap, xmap, p, smem, st); break;
        case T_Q8_0: rc = mmq_launch<T_Q8_0>(wmap, xmap, p, smem, st); break;
        case T_Q4_1: rc = mmq_launch<T_Q4_1>(wmap, xmap, p, smem, st); break;
        case T_Q5_0: rc = mmq_launch<T_Q5_0>(wmap, xmap, p, smem, st); break;
        case T_Q5_1: rc = mmq_launch<T_Q5_1>(wmap, xmap, p, smem, st); break;
        case T_Q4_K: rc = mmq_launch<T_Q4_K>(wmap, xmap, p, smem, st); break;
        case T_Q5_K: rc = mmq_launch<T_Q5_K>(wmap, xmap, p, smem, st); break;
        case T_Q6_K: rc = mmq_launch<T_Q6_K>(wmap, xmap, p, smem, st); break;
        case T_Q2_K: rc = mmq_launch<T_Q2_K>(wmap, xmap, p, smem, st); break;
        case T_Q3_K: rc = mmq_launch<T_Q3_K>(wmap, xmap, p, smem, st); break;
        case T_IQ4_NL: rc = mmq_launch<T_IQ4_NL>(wmap, xmap, p, smem, st); break;
        case T_IQ4_XS: rc = mmq_launch<T_IQ4_XS>(wmap, xmap, p, smem, st); break;
    }
    if (rc == 0 && p.ksplit > 1) rc = ksplit_reduce(p, st);
    return rc;
}

// Up to three weight matrices of one type, K and row stride that read the same activations (q/k/v of a prefill batch) in one
// grid: 8 + 8 + 32 row tiles fill the SMs where the k and v launches alone ran on 32 CTAs each.
extern "C" int b200_mul_mat_q_bf16_group(int type, int ng, const void * const * W, int64_t nb1, const void * Xb, float * const * dst,
                                         const int64_t * ds, const int * N, int K, int M, void * stream) {
    if (ng < 1 || ng > 3 || !W || !dst || !ds || !N) return -1;
    if (ng == 1) return b200_mul_mat_q_bf16(type, W[0], nb1, Xb, dst[0], ds[0], N[0], K, M, stream);
    if (M <= 0 || ((uintptr_t)Xb & 127)) return -1;
    int ntot = 0;
    for (int g = 0; g < ng; ++g) {
        if (!b200_mul_mat_q_supported(type, W[g], nb1, K) || N[g] <= 0 || !dst[g]) return -1;
        ntot += (N[g] + 127) / 128 * 128;
    }
    cudaStream_t st = (cudaStream_t)stream;
    mmq_params p = {};
    p.dst = dst[0]; p.ds = ds[0]; p.N = N[0]; p.M = M; p.K = K;
    p.tn = choose_tn(ntot, M);
    p.trace = nullptr;
    choose_rings(type, p);
    p.na = p.tn <= 128 ? 4 : NA_MAX;
    p.tmem_cols = p.tn <= 128 ? 256 : 512;
    if (mmq_smem(type, p.tn, p.nb, p.nr) > 232448) return -3;
    const size_t smem = mmq_smem(type, p.tn, p.nb, p.nr);
    p.ngroup = ng;
    CUtensorMap wmap[3], xmap;
    int y = 0;
    for (int g = 0; g < 3; ++g) {
        p.gy[g] = y;
        if (g >= ng) continue;
        p.gdst[g] = dst[g]; p.gds[g] = ds[g]; p.gN[g] = N[g];
        y += (N[g] + 127) / 128;
        const cuuint64_t dims[2] = {(cuuint64_t)((int64_t)(K / block_elems(type)) * block_bytes(type) / 2), (cuuint64_t)N[g]};
        const cuuint64_t strides[1] = {(cuuint64_t)nb1};
        const cuuint32_t box[2] = {(cuuint32_t)(raw_pitch(type) / 2), 128};
        const cuuint32_t es[2] = {1, 1};
        if (encode_tiled()(&wmap[g], CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, const_cast<void *>(W[g]), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) return -4;
    }
    p.gy[3] = y;
    for (int g = ng; g < 3; ++g) p.gy[g] = y;
    {
        const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)M};
        const cuuint64_t strides[1] = {(cuuint64_t)K * 2};
        const cuuint32_t box[2] = {64, (cuuint32_t)p.tn};
        const cuuint32_t es[2] = {1, 1};
        if (encode_tiled()(&xmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(Xb), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) return -5;
    }
    const CUtensorMap * w1 = &wmap[1], * w2 = ng > 2 ? &wmap[2] : &wmap[1];
    {   // one token tile: slices along K (columns of the slices: the groups side by side)
        int col = 0;
        for (int g = 0; g < 3; ++g) { p.gcol[g] = col; if (g < ng) col += N[g]; }
        p.part_ds = col;
        p.ksplit = 1;
        for (int g = 0; g < ng; ++g) {            // each matrix: the slice count of its own launch (8 slices of the whole width fit the buffer or nothing is split)
            float * pb = nullptr;
            p.gks[g] = plan_ksplit(type, K, M, p.tn, (N[g] + 127) / 128, N[g], st, &pb);
            if ((size_t)8 * M * p.part_ds * 4 > KSPLIT_BYTES) p.gks[g] = 1;
            if (p.gks[g] > 1) { p.part = pb; p.ksplit = std::max(p.ksplit, p.gks[g]); }
        }
        if (p.ksplit <= 1) { p.ksplit = 0; p.part = nullptr; }
    }
    int rc = -1;
    switch (type) {
        case T_Q4_0: rc = mmq_launch<T_Q4_0>(wmap[0], xmap, p, smem, st, 0, w1, w2); break;
        case T_Q8_0: rc = mmq_launch<T_Q8_0>(wmap[0], xmap, p, smem, st, 0, w1, w2); break;
        case T_Q4_1: rc = mmq_launch<T_Q4_1>(wmap[0], xmap, p, smem, st, 0, w1, w2); break;
        case T_Q5_0: rc = mmq_launch<T_Q5_0>(wmap[0], xmap, p, smem, st, 0, w1, w2); break;
        case T_Q5_1: rc = mmq_launch<T_Q5_1>(wmap[0], xmap, p, smem, st, 0, w1, w2); break;
        case T_Q4_K: rc = mmq_launch<T_Q4_K>(wmap[0], xmap, p, smem, st, 0, w1, w2); break;
        case T_Q5_K: rc = mmq_launch<T_Q5_K>(wmap[0], xmap, p, smem, st, 0, w1, w2); break;
        case T_Q6_K: rc = mmq_launch<T_Q6_K>(wmap[0], xmap, p, smem, st, 0, w1, w2); break;
        case T_Q2_K: rc = mmq_launch<T_Q2_K>(wmap[0], xmap, p, smem, st, 0, w1, w2); break;
        case T_Q3_K: rc = mmq_launch<T_Q3_K>(wmap[0], xmap, p, smem, st, 0, w1, w2); break;
        case T_IQ4_NL: rc = mmq_launch<T_IQ4_NL>(wmap[0], xmap, p, smem, st, 0, w1, w2); break;
        case T_IQ4_XS: rc = mmq_launch<T_IQ4_XS>(wmap[0], xmap, p, smem, st, 0, w1, w2); break;
    }
    if (rc == 0 && p.ksplit > 1) rc = ksplit_reduce(p, st);
    return rc;
}

extern "C" int b200_mul_mat_q(int type, const void * W, int64_t nb1, const float * X, int64_t xs, float * dst, int64_t ds,
                              int N, int K, int M, void * scratch, int64_t scratch_bytes, void * stream) {
    if (!b200_mul_mat_q_supported(type, W, nb1, K) || N <= 0 || M <= 0) return -1;
    if (!scratch || scratch_bytes < b200_mul_mat_q_scratch_bytes(K, M)) return -2;
    uint8_t * xb = (uint8_t *)(((uintptr_t)scratch + 127) & ~(uintptr_t)127);
    const int rc = b200_convert_acts_bf16(X, xs, xb, K, M, stream);
    if (rc) return rc;
    return b200_mul_mat_q_bf16(type, W, nb1, xb, dst, ds, N, K, M, stream);
}

// ---- MUL_MAT_ID for batches: rows sorted by expert on the device, one MMQ grid over (token tiles, row tiles, experts) ----
static int64_t align256(int64_t x) { return (x + 255) & ~(int64_t)255; }
extern "C" int64_t b200_mul_mat_id_q_scratch_bytes(int K, int n_used, int T, int n_expert) {
    const int64_t total = (int64_t)n_used * T;
    return align256(2 * (int64_t)n_expert * 4) + align256(total * 8) + align256(total * 8) + align256(total * K * 2 + 256 * (int64_t)K * 2) + 256;
}
// W: n_expert matrices [N][K] of one quantised type, nb1 bytes between rows, experts N * nb1 apart.  X: f32 activations of
// token t / slot u at X + t * x_sz + (u % x_ne1) * x_sy (floats).  ids: int32 [T][ids_s1], n_used per token.  dst row of
// (t, u) at dst + t * d_sz + u * d_sy (floats).  Returns -1 for shapes this path does not take (the caller falls back).
extern "C" int b200_mul_mat_id_q(int type, const void * W, int64_t nb1, int n_expert, const float * X, int64_t x_sy, int64_t x_sz, int x_ne1,
                                 const int32_t * ids, int64_t ids_s1, int n_used, int T, float * dst, int64_t d_sy, int64_t d_sz,
                                 int N, int K, void * scratch, int64_t scratch_bytes, void * stream) {
    if (!b200_mul_mat_q_supported(type, W, nb1, K) || N <= 0 || T <= 0 || n_used <= 0 || n_expert <= 0 || n_expert > 4096 || x_ne1 <= 0) return -1;
    if (!scratch || scratch_bytes < b200_mul_mat_id_q_scratch_bytes(K, n_used, T, n_expert)) return -2;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t total = (int64_t)n_used * T;
    uint8_t * sp = (uint8_t *)(((uintptr_t)scratch + 255) & ~(uintptr_t)255);
    int * cnt = (int *)sp; int * off = cnt + n_expert; sp += align256(2 * (int64_t)n_expert * 4);
    int2 * pairs = (int2 *)sp; sp += align256(total * 8);
    int64_t * row_off = (int64_t *)sp; sp += align256(total * 8);
    uint8_t * xb = sp;                                                 // bf16 [total (+ one tile of slack)][K]

    moe_sort_kernel<<<1, 1024, 2 * n_expert * sizeof(int), st>>>(ids, ids_s1, n_used, T, n_expert, cnt, off, pairs);
    B200_CUDA_CHECK(cudaGetLastError());
    moe_gather_kernel<<<(unsigned)total, 256, 0, st>>>(X, x_sy, x_sz, x_ne1, pairs, cnt, off, n_expert, (uint32_t *)xb, row_off, d_sy, d_sz, K);
    B200_CUDA_CHECK(cudaGetLastError());
    g_launches += 2;

    mmq_params p = {};
    p.dst = dst; p.ds = 0; p.N = N; p.K = K;
    p.M = (int)total;                                                  // worst case per expert: every row routes to it (CTAs past the real count leave at once)
    const int64_t avg = (total + n_expert - 1) / n_expert;
    p.tn = env_int("B200_MMQ_ID_TN", 0);
    if (p.tn != 32 && p.tn != 64 && p.tn != 128 && p.tn != 256) p.tn = avg <= 32 ? 32 : avg <= 64 ? 64 : avg <= 192 ? 128 : 256;
    p.trace = nullptr;
    choose_rings(type, p);
    p.na = p.tn <= 128 ? 4 : NA_MAX;
    p.tmem_cols = p.tn <= 128 ? 256 : 512;
    if (mmq_smem(type, p.tn, p.nb, p.nr) > 232448) return -3;
    const size_t smem = mmq_smem(type, p.tn, p.nb, p.nr);
    p.moe_cnt = cnt; p.moe_off = off; p.moe_row = row_off;

    CUtensorMap wmap, xmap;
    {   // the experts' matrices stacked: [n_expert * N][nb1 / 2] 16-bit words
        const cuuint64_t dims[2] = {(cuuint64_t)((int64_t)(K / block_elems(type)) * block_bytes(type) / 2), (cuuint64_t)N * (cuuint64_t)n_expert};
        const cuuint64_t strides[1] = {(cuuint64_t)nb1};
        const cuuint32_t box[2] = {(cuuint32_t)(raw_pitch(type) / 2), 128};
        const cuuint32_t es[2] = {1, 1};
        if (encode_tiled()(&wmap, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, const_cast<void *>(W), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) return -4;
    }
    {   // sorted bf16 rows [total][K]
        const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)total};
        const cuuint64_t strides[1] = {(cuuint64_t)K * 2};
        const cuuint32_t box[2] = {64, (cuuint32_t)p.tn};
        const cuuint32_t es[2] = {1, 1};
        if (encode_tiled()(&xmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, xb, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) return -5;
    }
    switch (type) {
        case T_Q4_0: return mmq_launch<T_Q4_0>(wmap, xmap, p, smem, st, n_expert);
        case T_Q8_0: return mmq_launch<T_Q8_0>(wmap, xmap, p, smem, st, n_expert);
        case T_Q4_1: return mmq_launch<T_Q4_1>(wmap, xmap, p, smem, st, n_expert);
        case T_Q5_0: return mmq_launch<T_Q5_0>(wmap, xmap, p, smem, st, n_expert);
        case T_Q5_1: return mmq_launch<T_Q5_1>(wmap, xmap, p, smem, st, n_expert);
        case T_Q4_K: return mmq_launch<T_Q4_K>(wmap, xmap, p, smem, st, n_expert);
        case T_Q5_K: return mmq_launch<T_Q5_K>(wmap, xmap, p, smem, st, n_expert);
        case T_Q6_K: return mmq_launch<T_Q6_K>(wmap, xmap, p, smem, st, n_expert);
        case T_Q2_K: return mmq_launch<T_Q2_K>(wmap, xmap, p, smem, st, n_expert);
        case T_Q3_K: return mmq_launch<T_Q3_K>(wmap, xmap, p, smem, st, n_expert);
        case T_IQ4_NL: return mmq_launch<T_IQ4_NL>(wmap, xmap, p, smem, st, n_expert);
        case T_IQ4_XS: return mmq_launch<T_IQ4_XS>(wmap, xmap, p, smem, st, n_expert);
    }
    return -1;
}
