/*
 * oracle.c -- CPU restatement of the reference's quantized-forward-pass arithmetic.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing under llamacog_b200/ may link, import or call this file;
 * only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg use it as a checker.
 *
 * Parity status: PINNED.  tests/test_oracle_golden.py checks every function below against
 *   (a) golden vectors in tests/golden/ produced by the reference's own libggml-base
 *       (tests/golden/make_golden.py, run in the build container against /root/reference), and
 *   (b) when oracle/_ref/libggml-ref.so is present, live against the reference functions.
 * Provenance of the goldens, vector by vector (tests/golden/make_golden.py docstring has the full list):
 *   reference-made   W_* deq_* R_* rdeq_* (ggml_quantize_chunk / dequantize_row_*), q80ref q81ref, h2f f2h: written by the
 *                    reference's own libggml-base compiled here;  tests/golden/q81_cuda_ref.npz: written by the reference's
 *                    own quantize_q8_1 CUDA kernel on a B200 (tests/golden/make_q81_cuda_golden.py)
 *   builder-derived  sumi_* / sumq (int32 block sums of vecdotq.cuh): no reference function exports them, so they are derived
 *                    by an independent numpy route from the reference's dequantised floats;  q81cuda: numpy restatement of
 *                    quantize.cu:4-49, confirmed equal to the reference kernel's bytes (Y512 in q81_cuda_ref.npz)
 *
 * Reference citations are relative to /root/reference/.
 * Plain C11, no SIMD, compile with -ffp-contract=off so that a*b-c is two roundings like the
 * reference's libggml-base (built without FMA).
 */
#include <stdint.h>
#include <string.h>
#include <math.h>
#include <stdlib.h>

#define QK 32     /* elements per q4_0 / q8_0 / q8_1 block   (ggml-common.h:166,208,216) */
#define QK_K 256  /* elements per K-quant super-block          (ggml-common.h:89) */

/* ---- ggml type ids (ggml/include/ggml.h:352-366) ---- */
enum { T_F32 = 0, T_F16 = 1, T_Q4_0 = 2, T_Q4_1 = 3, T_Q5_0 = 6, T_Q5_1 = 7, T_Q8_0 = 8, T_Q8_1 = 9, T_Q2_K = 10, T_Q3_K = 11, T_Q4_K = 12, T_Q5_K = 13, T_Q6_K = 14, T_IQ4_NL = 20, T_IQ4_XS = 23 };

/* ---- block layouts (ggml-common.h:167-172, 177-207 (q4_1 q5_0 q5_1), 209-214, 216-227, 285-296, 302-314, 320-326) ---- */
#pragma pack(push, 1)
typedef struct { uint16_t d; uint8_t qs[16]; } blk_q4_0;                                  /* 18 B */
typedef struct { uint16_t d; uint16_t m; uint8_t qs[16]; } blk_q4_1;                      /* 20 B */
typedef struct { uint16_t d; uint8_t qh[4]; uint8_t qs[16]; } blk_q5_0;                   /* 22 B */
typedef struct { uint16_t d; uint16_t m; uint8_t qh[4]; uint8_t qs[16]; } blk_q5_1;       /* 24 B */
typedef struct { uint16_t d; int8_t qs[32]; } blk_q8_0;                                   /* 34 B */
typedef struct { uint16_t d; uint16_t s; int8_t qs[32]; } blk_q8_1;                       /* 36 B */
typedef struct { uint8_t scales[16]; uint8_t qs[64]; uint16_t d; uint16_t dmin; } blk_q2_K;                 /* 84 B  (ggml-common.h:250-267) */
typedef struct { uint8_t hmask[32]; uint8_t qs[64]; uint8_t scales[12]; uint16_t d; } blk_q3_K;            /* 110 B (ggml-common.h:273-279) */
typedef struct { uint16_t d; uint16_t dmin; uint8_t scales[12]; uint8_t qs[128]; } blk_q4_K;              /* 144 B */
typedef struct { uint16_t d; uint8_t qs[16]; } blk_iq4_nl;                                                  /* 18 B  (ggml-common.h:405-410) */
typedef struct { uint16_t d; uint16_t scales_h; uint8_t scales_l[4]; uint8_t qs[128]; } blk_iq4_xs;         /* 136 B (ggml-common.h:412-418) */
typedef struct { uint16_t d; uint16_t dmin; uint8_t scales[12]; uint8_t qh[32]; uint8_t qs[128]; } blk_q5_K; /* 176 B */
typedef struct { uint8_t ql[128]; uint8_t qh[64]; int8_t scales[16]; uint16_t d; } blk_q6_K;              /* 210 B */
#pragma pack(pop)

int64_t oracle_block_bytes(int type) {
    switch (type) {
        case T_Q4_0: return 18; case T_Q8_0: return 34; case T_Q8_1: return 36;
        case T_Q4_1: return 20; case T_Q5_0: return 22; case T_Q5_1: return 24;
        case T_Q4_K: return 144; case T_Q5_K: return 176; case T_Q6_K: return 210; case T_Q2_K: return 84; case T_Q3_K: return 110;
        case T_IQ4_NL: return 18; case T_IQ4_XS: return 136;
        case T_F32: return 4; case T_F16: return 2;
    }
    return 0;
}
int64_t oracle_block_elems(int type) {
    switch (type) {
        case T_Q4_0: case T_Q8_0: case T_Q8_1: case T_Q4_1: case T_Q5_0: case T_Q5_1: case T_IQ4_NL: return 32;
        case T_Q4_K: case T_Q5_K: case T_Q6_K: case T_Q2_K: case T_Q3_K: case T_IQ4_XS: return 256;
        case T_F32: case T_F16: return 1;
    }
    return 0;
}

/* ---- IEEE binary16 <-> binary32, exact / round-to-nearest-even (ggml-impl.h fp16 helpers) ---- */
float oracle_fp16_to_fp32(uint16_t h) {
    uint32_t sign = (uint32_t)(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1f, man = h & 0x3ffu, bits;
    if (exp == 0) {
        if (man == 0) { bits = sign; }
        else { /* subnormal: normalise */
            int e = -1;
            do { e++; man <<= 1; } while ((man & 0x400u) == 0);
            bits = sign | ((uint32_t)(127 - 15 - e) << 23) | ((man & 0x3ffu) << 13);
        }
    } else if (exp == 31) { bits = sign | 0x7f800000u | (man << 13); }
    else { bits = sign | ((exp + 112) << 23) | (man << 13); }
    float f; memcpy(&f, &bits, 4); return f;
}

uint16_t oracle_fp32_to_fp16(float f) {
    uint32_t x; memcpy(&x, &f, 4);
    uint32_t sign = (x >> 16) & 0x8000u;
    uint32_t ax = x & 0x7fffffffu;
    if (ax >= 0x7f800000u) return (uint16_t)(sign | 0x7c00u | ((ax > 0x7f800000u) ? 0x200u : 0));
    if (ax >= 0x477ff000u) return (uint16_t)(sign | 0x7c00u);           /* rounds to inf */
    if (ax < 0x33000001u) return (uint16_t)sign;                          /* rounds to 0 (<= 2^-25) */
    int e = (int)(ax >> 23) - 127;
    uint32_t m = (ax & 0x7fffffu) | 0x800000u;
    int shift = (e < -14) ? (13 + (-14 - e)) : 13;                         /* bits to drop */
    uint32_t q = m >> shift, rem = m & ((1u << shift) - 1), half = 1u << (shift - 1);
    if (rem > half || (rem == half && (q & 1))) q++;
    uint32_t out = (e < -14) ? q : (((uint32_t)(e + 15) << 10) + (q - 0x400u));
    return (uint16_t)(sign | out);
}

/* 6-bit (scale,min) unpack of K-quants: ggml-quants.c:625-632 (get_scale_min_k4) */
static void scale_min_k4(int j, const uint8_t *q, int *sc, int *mn) {
    if (j < 4) { *sc = q[j] & 63; *mn = q[j + 4] & 63; }
    else { *sc = (q[j + 4] & 0x0F) | ((q[j - 4] >> 6) << 4); *mn = (q[j + 4] >> 4) | ((q[j] >> 6) << 4); }
}

/* ---- dequantize_row_*: ggml-quants.c:249-267 (q4_0), 343-357 (q8_0), 1274-1296 (q4_K),
 *      1480-1503 (q5_K), 1684-1713 (q6_K).  Must be bit-exact. ---- */
static void deq_q4_0(const blk_q4_0 *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK; i++) {
        float d = oracle_fp16_to_fp32(x[i].d);
        for (int j = 0; j < 16; j++) {
            y[i * QK + j]      = (float)((x[i].qs[j] & 0x0F) - 8) * d;
            y[i * QK + j + 16] = (float)((x[i].qs[j] >> 4) - 8) * d;
        }
    }
}
/* q4_1 (ggml-quants.c:269-288): nibble * d + m;  q5_0 (:290-314): (5-bit code - 16) * d, bit j / j + 16 of qh is the fifth bit of
 * element j / j + 16;  q5_1 (:316-341): 5-bit code * d + m.  Product and sum are two roundings (-ffp-contract=off). */
static uint32_t qh_word(const uint8_t *qh) { uint32_t v; memcpy(&v, qh, 4); return v; }
static void deq_q4_1(const blk_q4_1 *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK; i++) {
        float d = oracle_fp16_to_fp32(x[i].d), m = oracle_fp16_to_fp32(x[i].m);
        for (int j = 0; j < 16; j++) {
            y[i * QK + j]      = (float)(x[i].qs[j] & 0x0F) * d + m;
            y[i * QK + j + 16] = (float)(x[i].qs[j] >> 4) * d + m;
        }
    }
}
static void deq_q5_0(const blk_q5_0 *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK; i++) {
        float d = oracle_fp16_to_fp32(x[i].d);
        uint32_t qh = qh_word(x[i].qh);
        for (int j = 0; j < 16; j++) {
            y[i * QK + j]      = (float)((int)((x[i].qs[j] & 0x0F) | (((qh >> j) & 1) << 4)) - 16) * d;
            y[i * QK + j + 16] = (float)((int)((x[i].qs[j] >> 4) | (((qh >> (j + 16)) & 1) << 4)) - 16) * d;
        }
    }
}
static void deq_q5_1(const blk_q5_1 *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK; i++) {
        float d = oracle_fp16_to_fp32(x[i].d), m = oracle_fp16_to_fp32(x[i].m);
        uint32_t qh = qh_word(x[i].qh);
        for (int j = 0; j < 16; j++) {
            y[i * QK + j]      = (float)((x[i].qs[j] & 0x0F) | (((qh >> j) & 1) << 4)) * d + m;
            y[i * QK + j + 16] = (float)((x[i].qs[j] >> 4) | (((qh >> (j + 16)) & 1) << 4)) * d + m;
        }
    }
}
static void deq_q8_0(const blk_q8_0 *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK; i++) {
        float d = oracle_fp16_to_fp32(x[i].d);
        for (int j = 0; j < QK; j++) y[i * QK + j] = (float)x[i].qs[j] * d;
    }
}
static void deq_q4_K(const blk_q4_K *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        float d = oracle_fp16_to_fp32(x[i].d), mn = oracle_fp16_to_fp32(x[i].dmin);
        const uint8_t *q = x[i].qs;
        for (int p = 0; p < 4; p++) { /* pairs of 32-element sub-blocks share 32 bytes */
            int s0, m0, s1, m1;
            scale_min_k4(2 * p, x[i].scales, &s0, &m0);
            scale_min_k4(2 * p + 1, x[i].scales, &s1, &m1);
            float d0 = d * (float)s0, o0 = mn * (float)m0, d1 = d * (float)s1, o1 = mn * (float)m1;
            for (int l = 0; l < 32; l++) *y++ = d0 * (float)(q[l] & 0xF) - o0;
            for (int l = 0; l < 32; l++) *y++ = d1 * (float)(q[l] >> 4) - o1;
            q += 32;
        }
    }
}
static void deq_q5_K(const blk_q5_K *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        float d = oracle_fp16_to_fp32(x[i].d), mn = oracle_fp16_to_fp32(x[i].dmin);
        const uint8_t *ql = x[i].qs, *qh = x[i].qh;
        for (int p = 0; p < 4; p++) {
            int s0, m0, s1, m1;
            scale_min_k4(2 * p, x[i].scales, &s0, &m0);
            scale_min_k4(2 * p + 1, x[i].scales, &s1, &m1);
            float d0 = d * (float)s0, o0 = mn * (float)m0, d1 = d * (float)s1, o1 = mn * (float)m1;
            for (int l = 0; l < 32; l++) *y++ = d0 * (float)((ql[l] & 0xF) + (((qh[l] >> (2 * p)) & 1) ? 16 : 0)) - o0;
            for (int l = 0; l < 32; l++) *y++ = d1 * (float)((ql[l] >> 4) + (((qh[l] >> (2 * p + 1)) & 1) ? 16 : 0)) - o1;
            ql += 32;
        }
    }
}
/* integer value (already minus 32) of element e (0..255) of a q6_K block */
static int q6_val(const blk_q6_K *b, int e) {
    int h = e >> 7, r = e & 127, g = r >> 5, l = r & 31;       /* half, 32-group inside half, lane */
    uint8_t lo = b->ql[64 * h + 32 * (g & 1) + l];
    int nib = (g < 2) ? (lo & 0xF) : (lo >> 4);
    int hi = (b->qh[32 * h + l] >> (2 * g)) & 3;
    return (int)(int8_t)(nib | (hi << 4)) - 32;
}
static void deq_q6_K(const blk_q6_K *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        float d = oracle_fp16_to_fp32(x[i].d);
        for (int e = 0; e < QK_K; e++) {
            /* ggml-quants.c:1703-1706: y = d * sc * q, evaluated left to right in fp32 */
            float ds = d * (float)x[i].scales[e >> 4];
            y[i * QK_K + e] = ds * (float)q6_val(&x[i], e);
        }
    }
}

/* q2_K (ggml-quants.c:706-736): element e = 128 n + 32 j + l uses bits 2j..2j+1 of qs[32 n + l] and the 4-bit scale / 4-bit min pair
 * scales[8 n + 2 j + l / 16]:  y = (d * sc) * q - (dmin * mn), every product and the difference rounded on its own. */
static void deq_q2_K(const blk_q2_K *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        const float d = oracle_fp16_to_fp32(x[i].d), mn = oracle_fp16_to_fp32(x[i].dmin);
        for (int e = 0; e < QK_K; e++) {
            const int n = e >> 7, j = (e >> 5) & 3, l = e & 31;
            const uint8_t sc = x[i].scales[8 * n + 2 * j + (l >> 4)];
            const float dl = d * (float)(sc & 0xF), ml = mn * (float)(sc >> 4);
            y[i * QK_K + e] = dl * (float)((x[i].qs[32 * n + l] >> (2 * j)) & 3) - ml;
        }
    }
}
/* q3_K (ggml-quants.c:1050-1098): 6-bit scales (low nibbles in scales[0..7], high pairs in scales[8..11]) minus 32; code = 2 low bits of
 * qs minus 4 unless bit (4 n + j) of hmask[l] is set, -4..3;  y = (d * (sc - 32)) * code. */
static int q3_scale(const uint8_t *sc, int is) {
    const int b = is & 3, g = is >> 2;
    const int lo = g < 2 ? (sc[4 * g + b] & 0xF) : (sc[4 * (g - 2) + b] >> 4);
    const int hi = (sc[8 + b] >> (2 * g)) & 3;
    return (lo | (hi << 4)) - 32;
}
static int q3_code(const blk_q3_K *b, int e) {
    const int n = e >> 7, j = (e >> 5) & 3, l = e & 31;
    return (int)((b->qs[32 * n + l] >> (2 * j)) & 3) - ((b->hmask[l] & (1 << (4 * n + j))) ? 0 : 4);
}
static void deq_q3_K(const blk_q3_K *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        const float d = oracle_fp16_to_fp32(x[i].d);
        for (int e = 0; e < QK_K; e++) {
            const float dl = d * (float)q3_scale(x[i].scales, e >> 4);
            y[i * QK_K + e] = dl * (float)q3_code(&x[i], e);
        }
    }
}

/* iq4_nl / iq4_xs (ggml-quants.c:2428-2467): 4-bit indices into a fixed non-linear table (ggml-common.h:1077-1079); iq4_nl has one
 * f16 scale per 32 elements, iq4_xs a super-block scale times a 6-bit block scale minus 32.  Low nibbles are elements 0..15 of a
 * block, high nibbles 16..31. */
static const int8_t iq4_table[16] = { -127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113 };
static void deq_iq4_nl(const blk_iq4_nl *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK; i++) {
        const float d = oracle_fp16_to_fp32(x[i].d);
        for (int j = 0; j < 16; j++) {
            y[i * QK + j]      = d * (float)iq4_table[x[i].qs[j] & 0xF];
            y[i * QK + j + 16] = d * (float)iq4_table[x[i].qs[j] >> 4];
        }
    }
}
static int iq4_xs_scale(const blk_iq4_xs *b, int ib) {
    return (((b->scales_l[ib / 2] >> (4 * (ib % 2))) & 0xF) | (((b->scales_h >> (2 * ib)) & 3) << 4)) - 32;
}
static void deq_iq4_xs(const blk_iq4_xs *x, float *y, int64_t k) {
    for (int64_t i = 0; i < k / QK_K; i++) {
        const float d = oracle_fp16_to_fp32(x[i].d);
        for (int ib = 0; ib < 8; ib++) {
            const float dl = d * (float)iq4_xs_scale(&x[i], ib);
            for (int j = 0; j < 16; j++) {
                y[i * QK_K + 32 * ib + j]      = dl * (float)iq4_table[x[i].qs[16 * ib + j] & 0xF];
                y[i * QK_K + 32 * ib + j + 16] = dl * (float)iq4_table[x[i].qs[16 * ib + j] >> 4];
            }
        }
    }
}

int oracle_dequantize_row(int type, const void *x, float *y, int64_t k) {
    switch (type) {
        case T_Q4_0: deq_q4_0((const blk_q4_0 *)x, y, k); return 0;
        case T_Q8_0: deq_q8_0((const blk_q8_0 *)x, y, k); return 0;
        case T_Q4_1: deq_q4_1((const blk_q4_1 *)x, y, k); return 0;
        case T_Q5_0: deq_q5_0((const blk_q5_0 *)x, y, k); return 0;
        case T_Q5_1: deq_q5_1((const blk_q5_1 *)x, y, k); return 0;
        case T_Q4_K: deq_q4_K((const blk_q4_K *)x, y, k); return 0;
        case T_Q5_K: deq_q5_K((const blk_q5_K *)x, y, k); return 0;
        case T_Q6_K: deq_q6_K((const blk_q6_K *)x, y, k); return 0;
        case T_Q2_K: deq_q2_K((const blk_q2_K *)x, y, k); return 0;
        case T_Q3_K: deq_q3_K((const blk_q3_K *)x, y, k); return 0;
        case T_IQ4_NL: deq_iq4_nl((const blk_iq4_nl *)x, y, k); return 0;
        case T_IQ4_XS: deq_iq4_xs((const blk_iq4_xs *)x, y, k); return 0;
        case T_F32: memcpy(y, x, (size_t)k * 4); return 0;
        case T_F16: for (int64_t i = 0; i < k; i++) y[i] = oracle_fp16_to_fp32(((const uint16_t *)x)[i]); return 0;
    }
    return -1;
}

/* ---- activation quantisation ------------------------------------------------------------- */

/* CUDA-style q8_1 (ggml-cuda/quantize.cu:4-49): d = amax/127 (a true division), q = roundf(x/d)
 * (0 when amax == 0), s = sum of the UNQUANTISED x, reduced as a 32-lane xor-butterfly
 * (offsets 16,8,4,2,1; common.cuh:350-356), both stored as binary16.  k is padded with zeros up
 * to k_pad (multiple of 32; the reference pads rows to 512, mmvq.cu:527-528). */
void oracle_quantize_row_q8_1_cuda(const float *x, void *vy, int64_t k, int64_t k_pad) {
    blk_q8_1 *y = (blk_q8_1 *)vy;
    for (int64_t ib = 0; ib < k_pad / QK; ib++) {
        float v[QK], s[QK], amax = 0.0f;
        for (int j = 0; j < QK; j++) {
            int64_t i = ib * QK + j;
            v[j] = i < k ? x[i] : 0.0f;
            s[j] = v[j];
            amax = fmaxf(amax, fabsf(v[j]));
        }
        for (int off = 16; off > 0; off >>= 1) {
            float t[QK];
            for (int j = 0; j < QK; j++) t[j] = s[j] + s[j ^ off];
            memcpy(s, t, sizeof s);
        }
        const float d = amax / 127;
        for (int j = 0; j < QK; j++) y[ib].qs[j] = amax == 0.0f ? 0 : (int8_t)roundf(v[j] / d);
        y[ib].d = oracle_fp32_to_fp16(d);
        y[ib].s = oracle_fp32_to_fp16(s[0]);
    }
}

/* CPU-style reference quantisers (ggml-quants.c:187-210 q8_0_ref, 213-247 q8_1_ref): multiply by
 * the reciprocal id = 1/d.  q8_0_ref is what the f32->q8_0 KV-cache store must reproduce. */
void oracle_quantize_row_q8_0_ref(const float *x, void *vy, int64_t k) {
    blk_q8_0 *y = (blk_q8_0 *)vy;
    for (int64_t ib = 0; ib < k / QK; ib++) {
        float amax = 0.0f;
        for (int j = 0; j < QK; j++) amax = fmaxf(amax, fabsf(x[ib * QK + j]));
        const float d = amax / 127, id = d ? 1.0f / d : 0.0f;
        y[ib].d = oracle_fp32_to_fp16(d);
        for (int j = 0; j < QK; j++) y[ib].qs[j] = (int8_t)roundf(x[ib * QK + j] * id);
    }
}
void oracle_quantize_row_q8_1_ref(const float *x, void *vy, int64_t k) {
    blk_q8_1 *y = (blk_q8_1 *)vy;
    for (int64_t ib = 0; ib < k / QK; ib++) {
        float amax = 0.0f;
        for (int j = 0; j < QK; j++) amax = fmaxf(amax, fabsf(x[ib * QK + j]));
        const float d = amax / 127, id = d ? 1.0f / d : 0.0f;
        int sum = 0;
        for (int j = 0; j < QK; j++) { y[ib].qs[j] = (int8_t)roundf(x[ib * QK + j] * id); sum += y[ib].qs[j]; }
        y[ib].d = oracle_fp32_to_fp16(d);
        y[ib].s = oracle_fp32_to_fp16((float)sum * d);
    }
}

/* ---- integer weight codes ---------------------------------------------------------------- */
/* The unsigned/offset integer code w[e] each format stores for element e, such that the
 * dequantised value is  scale(e) * w[e] - offset(e).  These are the operands of the int32 dot
 * products in ggml-cuda/vecdotq.cuh (q4_0: nibble 0..15 with the -8 folded into the s8 term,
 * :25-44; q8_0: int8, :153-165; q4_K: nibble 0..15, :359-381; q5_K: 0..31, :415-444; q6_K: -32..31
 * after __vsubss4, :478-498). */
static void int_codes(int type, const void *row, int64_t k, int *w) {
    if (type == T_Q4_0) {
        const blk_q4_0 *b = (const blk_q4_0 *)row;
        for (int64_t i = 0; i < k / QK; i++) for (int j = 0; j < 16; j++) {
            w[i * QK + j] = b[i].qs[j] & 0xF; w[i * QK + j + 16] = b[i].qs[j] >> 4;
        }
    } else if (type == T_Q4_1) {      /* nibble 0..15 (vecdotq.cuh:49-80) */
        const blk_q4_1 *b = (const blk_q4_1 *)row;
        for (int64_t i = 0; i < k / QK; i++) for (int j = 0; j < 16; j++) {
            w[i * QK + j] = b[i].qs[j] & 0xF; w[i * QK + j + 16] = b[i].qs[j] >> 4;
        }
    } else if (type == T_Q5_0 || type == T_Q5_1) {      /* 5-bit code 0..31, the -16 of q5_0 folded into the s8 term (vecdotq.cuh:82-150) */
        for (int64_t i = 0; i < k / QK; i++) {
            const uint8_t *blk = (const uint8_t *)row + i * (type == T_Q5_0 ? 22 : 24);
            const uint8_t *qs = blk + (type == T_Q5_0 ? 6 : 8);
            const uint32_t qh = qh_word(blk + (type == T_Q5_0 ? 2 : 4));
            for (int j = 0; j < 16; j++) {
                w[i * QK + j]      = (int)((qs[j] & 0xF) | (((qh >> j) & 1) << 4));
                w[i * QK + j + 16] = (int)((qs[j] >> 4)  | (((qh >> (j + 16)) & 1) << 4));
            }
        }
    } else if (type == T_Q8_0) {
        const blk_q8_0 *b = (const blk_q8_0 *)row;
        for (int64_t i = 0; i < k / QK; i++) for (int j = 0; j < QK; j++) w[i * QK + j] = b[i].qs[j];
    } else if (type == T_Q4_K) {
        const blk_q4_K *b = (const blk_q4_K *)row;
        for (int64_t i = 0; i < k / QK_K; i++) for (int p = 0; p < 4; p++) for (int l = 0; l < 32; l++) {
            w[i * QK_K + 64 * p + l] = b[i].qs[32 * p + l] & 0xF; w[i * QK_K + 64 * p + 32 + l] = b[i].qs[32 * p + l] >> 4;
        }
    } else if (type == T_Q5_K) {
        const blk_q5_K *b = (const blk_q5_K *)row;
        for (int64_t i = 0; i < k / QK_K; i++) for (int p = 0; p < 4; p++) for (int l = 0; l < 32; l++) {
            w[i * QK_K + 64 * p + l]      = (b[i].qs[32 * p + l] & 0xF) | (((b[i].qh[l] >> (2 * p)) & 1) << 4);
            w[i * QK_K + 64 * p + 32 + l] = (b[i].qs[32 * p + l] >> 4)  | (((b[i].qh[l] >> (2 * p + 1)) & 1) << 4);
        }
    } else if (type == T_Q6_K) {
        const blk_q6_K *b = (const blk_q6_K *)row;
        for (int64_t i = 0; i < k / QK_K; i++) for (int e = 0; e < QK_K; e++) w[i * QK_K + e] = q6_val(&b[i], e);
    } else if (type == T_Q2_K) {      /* 0..3 (vecdotq.cuh:218-243) */
        const blk_q2_K *b = (const blk_q2_K *)row;
        for (int64_t i = 0; i < k / QK_K; i++) for (int e = 0; e < QK_K; e++)
            w[i * QK_K + e] = (b[i].qs[32 * (e >> 7) + (e & 31)] >> (2 * ((e >> 5) & 3))) & 3;
    } else if (type == T_IQ4_NL) {    /* table values (vecdotq.cuh: vec_dot_iq4_nl_q8_1) */
        const blk_iq4_nl *b = (const blk_iq4_nl *)row;
        for (int64_t i = 0; i < k / QK; i++) for (int j = 0; j < 16; j++) {
            w[i * QK + j] = iq4_table[b[i].qs[j] & 0xF]; w[i * QK + j + 16] = iq4_table[b[i].qs[j] >> 4];
        }
    } else if (type == T_IQ4_XS) {
        const blk_iq4_xs *b = (const blk_iq4_xs *)row;
        for (int64_t i = 0; i < k / QK_K; i++) for (int ib = 0; ib < 8; ib++) for (int j = 0; j < 16; j++) {
            w[i * QK_K + 32 * ib + j] = iq4_table[b[i].qs[16 * ib + j] & 0xF]; w[i * QK_K + 32 * ib + j + 16] = iq4_table[b[i].qs[16 * ib + j] >> 4];
        }
    } else if (type == T_Q3_K) {      /* -4..3 after __vsubss4 (vecdotq.cuh:301-331) */
        const blk_q3_K *b = (const blk_q3_K *)row;
        for (int64_t i = 0; i < k / QK_K; i++) for (int e = 0; e < QK_K; e++) w[i * QK_K + e] = q3_code(&b[i], e);
    }
}

/* Per 32-element sub-block of one weight row against one q8_1 activation row:
 *   sumi[b] = sum_j w[32b+j] * q[32b+j]   (the dp4a accumulations of vecdotq.cuh), int32, EXACT
 *   sumq[b] = sum_j q[32b+j]              (dot2 of vecdotq.cuh:376,437), int32, EXACT
 * Integer sums are order-independent, so any thread layout must reproduce them bit for bit. */
int oracle_block_sums(int type, const void *wrow, const void *q8_1, int64_t k, int32_t *sumi, int32_t *sumq) {
    if (oracle_block_elems(type) == 0 || k % oracle_block_elems(type)) return -1;
    int *w = (int *)malloc(sizeof(int) * (size_t)k);
    int_codes(type, wrow, k, w);
    const blk_q8_1 *y = (const blk_q8_1 *)q8_1;
    for (int64_t b = 0; b < k / QK; b++) {
        int32_t si = 0, sq = 0;
        for (int j = 0; j < QK; j++) { si += w[b * QK + j] * y[b].qs[j]; sq += y[b].qs[j]; }
        sumi[b] = si; sumq[b] = sq;
    }
    free(w);
    return 0;
}

/* ---- quantised GEMV / skinny GEMM reference ------------------------------------------------
 * dst[n, m] = sum_k dequant(W[n, k]) * dequant_q8_1(X[m, k]) with X quantised CUDA-style, i.e. the
 * quantity ggml-cuda's mul_mat_vec_q (mmvq.cu:140-224) approximates in fp32; here accumulated in
 * double from the exact integer sums so the GPU result can be checked to ~1e-6 relative.
 * For q4_0 the reference subtracts 8*s8 with s8 = fp16(sum x) (vecdotq.cuh:43); use_s8 selects
 * that literal formula, otherwise the exact 8*d8*sumq is used (what our kernel computes). */
int oracle_mul_mat_vec_q(int type, const void *W, int64_t row_stride_bytes, const float *X, int64_t x_stride,
                         float *dst, int64_t dst_stride, int64_t N, int64_t K, int64_t M, int use_s8) {
    const int64_t be = oracle_block_elems(type);
    if (be == 0 || K % be) return -1;
    const int64_t nb32 = K / QK;
    blk_q8_1 *y = (blk_q8_1 *)malloc((size_t)nb32 * sizeof(blk_q8_1));
    int32_t *si = (int32_t *)malloc((size_t)nb32 * 4), *sq = (int32_t *)malloc((size_t)nb32 * 4);
    for (int64_t m = 0; m < M; m++) {
        oracle_quantize_row_q8_1_cuda(X + m * x_stride, y, K, K);
        for (int64_t n = 0; n < N; n++) {
            const uint8_t *row = (const uint8_t *)W + n * row_stride_bytes;
            oracle_block_sums(type, row, y, K, si, sq);
            double acc = 0.0;
            for (int64_t b = 0; b < nb32; b++) {
                const double d8 = oracle_fp16_to_fp32(y[b].d), s8 = oracle_fp16_to_fp32(y[b].s);
                if (type == T_Q4_0) {
                    const double d = oracle_fp16_to_fp32(((const blk_q4_0 *)row)[b].d);
                    acc += use_s8 ? d * (si[b] * d8 - 8.0 * s8) : d * d8 * (si[b] - 8.0 * sq[b]);
                } else if (type == T_Q4_1 || type == T_Q5_1) {
                    /* d * d8 * sumi + m * s8 (vecdotq.cuh:76-79, 146-149): s8 = fp16(sum x) in the reference; exact m * d8 * sumq here */
                    const uint8_t *blk = row + b * (type == T_Q4_1 ? 20 : 24);
                    const double d = oracle_fp16_to_fp32(((const uint16_t *)blk)[0]), mm = oracle_fp16_to_fp32(((const uint16_t *)blk)[1]);
                    acc += use_s8 ? d * d8 * si[b] + mm * s8 : d * d8 * si[b] + mm * d8 * sq[b];
                } else if (type == T_Q5_0) {
                    const double d = oracle_fp16_to_fp32(((const blk_q5_0 *)row)[b].d);
                    acc += use_s8 ? d * (si[b] * d8 - 16.0 * s8) : d * d8 * (si[b] - 16.0 * sq[b]);
                } else if (type == T_IQ4_NL) {
                    acc += (double)oracle_fp16_to_fp32(((const blk_iq4_nl *)row)[b].d) * d8 * si[b];
                } else if (type == T_IQ4_XS) {
                    const blk_iq4_xs *blk = (const blk_iq4_xs *)row + b / 8;
                    acc += (double)oracle_fp16_to_fp32(blk->d) * iq4_xs_scale(blk, (int)(b % 8)) * d8 * si[b];
                } else if (type == T_Q8_0) {
                    acc += (double)oracle_fp16_to_fp32(((const blk_q8_0 *)row)[b].d) * d8 * si[b];
                } else if (type == T_Q4_K || type == T_Q5_K) {
                    const uint8_t *blk = row + (b / 8) * (type == T_Q4_K ? 144 : 176);
                    const double d = oracle_fp16_to_fp32(((const uint16_t *)blk)[0]), dm = oracle_fp16_to_fp32(((const uint16_t *)blk)[1]);
                    int sc, mn; scale_min_k4((int)(b % 8), blk + 4, &sc, &mn);
                    acc += d * d8 * ((double)si[b] * sc) - dm * d8 * ((double)sq[b] * mn);
                } else if (type == T_Q2_K || type == T_Q3_K) {
                    /* two 16-element scale groups per 32 sub-block, as q6_K */
                    for (int g = 0; g < 2; g++) {
                        int32_t sw = 0, sx = 0;
                        const int e0 = (int)(b % 8) * 32 + g * 16;
                        if (type == T_Q2_K) {
                            const blk_q2_K *blk = (const blk_q2_K *)row + b / 8;
                            for (int j = 0; j < 16; j++) { const int e = e0 + j; sw += (int)((blk->qs[32 * (e >> 7) + (e & 31)] >> (2 * ((e >> 5) & 3))) & 3) * y[b].qs[g * 16 + j]; sx += y[b].qs[g * 16 + j]; }
                            const uint8_t sc = blk->scales[e0 >> 4];
                            acc += (double)oracle_fp16_to_fp32(blk->d) * d8 * ((double)sw * (sc & 0xF)) - (double)oracle_fp16_to_fp32(blk->dmin) * d8 * ((double)sx * (sc >> 4));
                        } else {
                            const blk_q3_K *blk = (const blk_q3_K *)row + b / 8;
                            for (int j = 0; j < 16; j++) sw += q3_code(blk, e0 + j) * y[b].qs[g * 16 + j];
                            acc += (double)oracle_fp16_to_fp32(blk->d) * d8 * ((double)sw * q3_scale(blk->scales, e0 >> 4));
                        }
                    }
                } else if (type == T_Q6_K) {
                    /* two 16-element scale groups per 32 sub-block: redo the split */
                    const blk_q6_K *blk = (const blk_q6_K *)row + b / 8;
                    const double d = oracle_fp16_to_fp32(blk->d);
                    for (int g = 0; g < 2; g++) {
                        int32_t s = 0;
                        for (int j = 0; j < 16; j++) { int e = (int)(b % 8) * 32 + g * 16 + j; s += q6_val(blk, e) * y[b].qs[g * 16 + j]; }
                        acc += d * d8 * ((double)s * blk->scales[(b % 8) * 2 + g]);
                    }
                }
            }
            dst[m * dst_stride + n] = (float)acc;
        }
    }
    free(y); free(si); free(sq);
    return 0;
}

/* plain dequantise-then-fp32-dot GEMV (no activation quantisation): the quantity the CPU backend
 * approximates; used to state NMSE the way tests/test-backend-ops.cpp:177-190 does. */
int oracle_mul_mat_f32(int type, const void *W, int64_t row_stride_bytes, const float *X, int64_t x_stride,
                       float *dst, int64_t dst_stride, int64_t N, int64_t K, int64_t M) {
    float *wf = (float *)malloc((size_t)K * 4);
    for (int64_t n = 0; n < N; n++) {
        if (oracle_dequantize_row(type, (const uint8_t *)W + n * row_stride_bytes, wf, K)) { free(wf); return -1; }
        for (int64_t m = 0; m < M; m++) {
            double acc = 0.0;
            for (int64_t k = 0; k < K; k++) acc += (double)wf[k] * X[m * x_stride + k];
            dst[m * dst_stride + n] = (float)acc;
        }
    }
    free(wf);
    return 0;
}
