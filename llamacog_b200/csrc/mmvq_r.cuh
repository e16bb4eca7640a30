// mmvq_r.cuh -- batch-1 decode GEMV with REGISTER-RESIDENT activations (sm_100a), v3 of the decode mat-vec.
//
// Same operator as mmvq.cuh (reference ggml-cuda/mmvq.cu:140-224 mul_mat_vec_q + quantize.cu:4-49), for the case
// that dominates a decode step: one activation column (M = 1), already quantised by the kernel that produced it
// (b200_quantize_acts layout, mmvq_units.cuh), K-quant weights in whole contiguous rows.
//
// Why a second kernel: in mmvq.cuh every thread reads its 256-weight item (144 B) AND the matching activations
// (400 B) from shared memory, so a 14336x4096 Q4_K GEMV moves 5x more bytes through the 128 B/clk shared-memory
// pipe than through HBM and is bound by it (ncu: 0.5 IPC per scheduler, r01_mmvq_v2_notes.md).  Here a thread owns
// a fixed HALF-ITEM column h of the row (128 weights: blocks 4h..4h+3 of its super-block) for the whole life of
// the CTA, so its 128 int8 activations + 4 scales + 4 sums live in ~40 registers, loaded ONCE (one bulk copy of the column into shared memory per CTA, then 12 LDS.128 per thread); per row it
// reads only 16 B of header + 64 B of codes from shared memory.  Shared-memory traffic drops from 688 to 304 B per
// 256 weights, instructions from ~300 to ~2 x 120.
//
//   warp 0      producer + epilogue: 1-D TMA bulk copies of R whole rows per stage (issued before the PDL wait,
//               so the weights stream while the previous kernel of the decode step is still running), then per
//               stage the fixed-order sum of the per-warp partials and a coalesced store of dst.
//   warps 1..8  consumers: thread (rsub, h) with h = ct % TPR, rsub = ct / TPR (TPR = threads per row = the power
//               of two >= 2*K/256; lanes h >= 2*K/256 idle).  A pass covers 256/TPR rows; two passes are in
//               flight per thread (independent shared loads and dp4a chains), row sums by xor-shuffle over
//               min(TPR, 32) lanes + one partial per warp in shared memory.
//
// The CTA is sized for HALF an SM (<= 113 KB, 288 threads, <= 112 registers) and the default grid is one CTA per
// SM: two consecutive GEMVs of a decode step are co-resident, the second one prefetching its weights into its own
// stages while the first computes -- HBM does not idle at kernel boundaries.
//
// Integer block sums are those of vecdotq.cuh:359-381 (q4_K) and 478-498 (q6_K), bit-exact (dbg hook); the fp32
// combination order is fixed (deterministic) but ours.
#pragma once
#include "mmvq_units.cuh"
#include "rope_units.cuh"

namespace b200 {

struct mmvqr_params {
    const uint8_t * Wg[3];   // weight matrices (contiguous 16-byte aligned rows of rb bytes)
    float * Dg[3];           // outputs
    int cta0[4];             // CTAs [cta0[g], cta0[g+1]) work on matrix g
    int rows_qg[3], rows_remg[3];
    int ngroup;
    const uint8_t * Xq;      // quantised activations of the one column
    int K, BPR;              // BPR = K/256 items per row
    int TPR, lgTPR;          // threads per row (power of two >= 2*BPR, <= 256)
    int rb;                  // bytes per row
    int R;                   // rows per stage (multiple of 256/TPR)
    int r0, r1, r2;          // rows of the first three stages of a CTA (ramp: <= R)
    int stagger;             // tuning: stages 1.. are requested only when stage 0 has landed
    int nstages, stage_bytes;
    int32_t * dbg_sumi;      // parity hook: [N][K/32] int32 block sums (single matrix only)
    // fused RMS_NORM + MUL prologue (NORM kernels): the activations are rms_norm(nx) * nw, computed by every CTA for itself;
    // CTA 0 also writes the two tensors of the original nodes and the quantised column (for later mat-muls that read it)
    const float * nx; const float * nw; float * nnorm; float * nout; uint8_t * nxq; float neps;
    // fused residual ADD epilogue (single matrix): dsum[n] = dst[n] + resid[n]
    const float * resid; float * dsum; int resid_rows;
    // fused ROPE + KV-cache store epilogue of a q/k/v launch (b200_gemv_rope): per matrix mode 0 none, 1 out = rope(row sums),
    // 2 also cell = convert(out), 3 cell = convert(row sums).  Row ranges of the CTAs are even-aligned (pairs stay together).
    int rope_on, rope_rows;
    int rmode[3]; float * rout[3]; char * rcell[3]; char * const * rslot[3]; int rctype[3], rcesz[3], rhd[3];
    const int32_t * rpos; const float * rff; rope_dev rp;
    hs_params hs;            // row-split hand-over folded into the launch (common.cuh)
    // MUL_MAT_ID at batch 1 (reference mmvq.cu:161-164 ids_dst path): matrix g is expert ids[g] of the stack at Wg[g] (w_expert bytes per
    // expert), read on the device after the PDL wait (the ids come out of the router kernels of the same step); its activations
    // are column xq_col[g] of Xq
    const int32_t * ids; int64_t w_expert; int xq_col[3];
    // gate (matrix 0) / up (matrix 1) launch in PAIR mode: every CTA owns the same whole 32-row blocks of BOTH matrices (it streams its
    // gate rows, then its up rows: 2 * nst stages), keeps its gate sums in shared memory and, as the up sums arrive, stores
    // silu(gate), silu(gate) * up and finally the q8 activations of its blocks (the arithmetic of unary_mul_kernel<SILU>, fused.cu).
    // No cross-CTA traffic.  (The first version dealt gate and up rows to different CTAs and counted 32-row blocks with atomics:
    // two fences, an atomic and a dependent load on the tail of every CTA cost what the removed launch had cost.)
    float * sm_act; float * sm_out; uint8_t * sm_xq;
    // pair mode over the chosen experts of a MoE layer (ids != nullptr): group g = slot g; Wg[g] is the gate stack and Wup the up stack
    // (the expert offset ids[g] * w_expert goes on both), Dg[g] / Du[g] the gate / up results of the slot, sm_act / sm_out advance by
    // sm_slot floats and the q8 column by sm_xq_col bytes per slot
    const uint8_t * Wup; float * Du[3]; int64_t sm_slot, sm_xq_col;
};

struct hacts { int q[4][8]; float d8[4], dsq[4]; int m32a[4], m32b[4]; };

// sp: shared address of the thread's 192 bytes (4 blocks of 48) inside the staged copy of Xq
__device__ __forceinline__ void load_hacts(uint32_t sp, hacts & a) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int4 t0 = lds128(sp + 48 * k), t1 = lds128(sp + 48 * k + 16), t2 = lds128(sp + 48 * k + 32);
        a.q[k][0] = t0.x; a.q[k][1] = t0.y; a.q[k][2] = t0.z; a.q[k][3] = t0.w;
        a.q[k][4] = t1.x; a.q[k][5] = t1.y; a.q[k][6] = t1.z; a.q[k][7] = t1.w;
        a.d8[k] = __int_as_float(t2.x); a.dsq[k] = __int_as_float(t2.y); a.m32a[k] = t2.z; a.m32b[k] = t2.w;
    }
}

// raw half-item of weights in registers
template <int TYPE> struct hweights;
template <> struct hweights<T_Q4_K> { int hd[4]; int w[16]; };
template <> struct hweights<T_Q6_K> { int ql[16]; int qh[8]; int sc[2]; uint32_t d; };
template <> struct hweights<T_Q8_0> { int w[34]; };      // four blocks of {d (f16), qs int8[32]}: 136 bytes, 8-byte aligned
template <> struct hweights<T_Q4_0> { int w[18]; };      // four blocks of {d (f16), qs[16]}: 72 bytes, 8-byte aligned
template <> struct hweights<T_Q5_K> { int hd[4]; int qh[8]; int w[16]; };      // header, the fifth bits of all 256 weights, this half's 64 code bytes
// the formats of the round-2 widening (a half item = four 32-blocks, or half a super-block):
template <> struct hweights<T_Q4_1> { int w[20]; };      // 4 x {d, m, qs[16]}: 80 bytes, 16-byte aligned
template <> struct hweights<T_Q5_0> { int w[22]; };      // 4 x {d, qh[4], qs[16]}: 88 bytes, 8-byte aligned
template <> struct hweights<T_Q5_1> { int w[24]; };      // 4 x {d, m, qh[4], qs[16]}: 96 bytes, 16-byte aligned
template <> struct hweights<T_IQ4_NL> { int w[18]; };    // as Q4_0
template <> struct hweights<T_Q2_K> { int sc[2]; int q[8]; int dm; };                 // this half's 8 scale bytes and 32 code bytes, d | dmin
template <> struct hweights<T_Q3_K> { int hm[8]; int q[8]; int sc[3]; uint32_t d; };   // hmask (all of it), this half's 32 code bytes, the 12 scale bytes, d
template <> struct hweights<T_IQ4_XS> { int hd[2]; int q[16]; };                       // d | scales_h, scales_l; this half's 64 code bytes

// ip: shared address of the item, hh: which half
template <int TYPE> __device__ __forceinline__ void load_hweights(uint32_t ip, int hh, hweights<TYPE> & w) {
    if constexpr (TYPE == T_Q4_K) {
        lds_words<4, true>(ip, w.hd);
        lds_words<16, true>(ip + 16 + 64 * hh, w.w);
    } else if constexpr (TYPE == T_Q8_0 || TYPE == T_Q4_0 || TYPE == T_IQ4_NL || TYPE == T_Q4_1 || TYPE == T_Q5_0 || TYPE == T_Q5_1) {
        constexpr int HB = TYPE == T_Q8_0 ? 136 : TYPE == T_Q4_1 ? 80 : TYPE == T_Q5_0 ? 88 : TYPE == T_Q5_1 ? 96 : 72;      // 8-byte aligned halves
#pragma unroll
        for (int i = 0; i < HB / 8; ++i) {
            uint32_t lo, hi;
            asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(lo), "=r"(hi) : "r"(ip + HB * hh + 8 * i));
            w.w[2 * i] = (int)lo; w.w[2 * i + 1] = (int)hi;
        }
    } else if constexpr (TYPE == T_Q5_K) {
        lds_words<4, true>(ip, w.hd);
        lds_words<8, true>(ip + 16, w.qh);
        lds_words<16, true>(ip + 48 + 64 * hh, w.w);
    } else if constexpr (TYPE == T_Q2_K) {         // 84-byte items: 4-byte aligned
#pragma unroll
        for (int i = 0; i < 2; ++i) w.sc[i] = (int)lds32(ip + 8 * hh + 4 * i);
#pragma unroll
        for (int i = 0; i < 8; ++i) w.q[i] = (int)lds32(ip + 16 + 32 * hh + 4 * i);
        w.dm = (int)lds32(ip + 80);
    } else if constexpr (TYPE == T_Q3_K) {         // 110-byte items: 2-byte aligned
        lds_words_a2<8>(ip, w.hm);
        lds_words_a2<8>(ip + 32 + 32 * hh, w.q);
        lds_words_a2<3>(ip + 96, w.sc);
        w.d = lds16(ip + 108);
    } else if constexpr (TYPE == T_IQ4_XS) {       // 136-byte items: 8-byte aligned
#pragma unroll
        for (int i = 0; i < 2; ++i) w.hd[i] = (int)lds32(ip + 4 * i);
#pragma unroll
        for (int i = 0; i < 16; ++i) w.q[i] = (int)lds32(ip + 8 + 64 * hh + 4 * i);
    } else {
        lds_words_a2<16>(ip + 64 * hh, w.ql);
        lds_words_a2<8>(ip + 128 + 32 * hh, w.qh);
        lds_words_a2<2>(ip + 192 + 8 * hh, w.sc);
        w.d = lds16(ip + 208);
    }
}

template <int TYPE, bool DBG>
__device__ __forceinline__ float half_dot(const hweights<TYPE> & w, const hacts & a, int hh, int32_t * dbg) {
    if constexpr (TYPE == T_Q4_K) {
        const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&w.hd[0]));
        const uint32_t h1 = (uint32_t)w.hd[1], h2 = (uint32_t)w.hd[2], h3 = (uint32_t)w.hd[3];
        // 6-bit (scale, min) of blocks 4hh..4hh+3 as packed bytes (reference ggml-quants.c:625-632)
        const uint32_t sc4 = hh ? ((h3 & 0x0F0F0F0Fu) | ((h1 >> 2) & 0x30303030u)) : (h1 & 0x3F3F3F3Fu);
        const uint32_t mn4 = hh ? (((h3 >> 4) & 0x0F0F0F0Fu) | ((h2 >> 2) & 0x30303030u)) : (h2 & 0x3F3F3F3Fu);
        float accd = 0.0f, accm = 0.0f;
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            int s0a = 0, s0b = 0, s1a = 0, s1b = 0;
#pragma unroll
            for (int j = 0; j < 8; j += 2) {
                const uint32_t q0 = (uint32_t)w.w[8 * p + j], q1 = (uint32_t)w.w[8 * p + j + 1];
                s0a = dp4a_us(q0 & 0x0F0F0F0Fu, a.q[2 * p][j], s0a);
                s0b = dp4a_us(q1 & 0x0F0F0F0Fu, a.q[2 * p][j + 1], s0b);
                s1a = dp4a_us(q0 & 0xF0F0F0F0u, a.q[2 * p + 1][j], s1a);       // 16 x nibble, divided out below
                s1b = dp4a_us(q1 & 0xF0F0F0F0u, a.q[2 * p + 1][j + 1], s1b);
            }
            const int s0 = s0a + s0b, s1 = (s1a + s1b) >> 4;
            if constexpr (DBG) { dbg[2 * p] = s0; dbg[2 * p + 1] = s1; }
            accd = fmaf(a.d8[2 * p], i2f_small(s0) * byte2f(sc4, 2 * p), accd);
            accd = fmaf(a.d8[2 * p + 1], i2f_small(s1) * byte2f(sc4, 2 * p + 1), accd);
            accm = fmaf(a.dsq[2 * p], byte2f(mn4, 2 * p), accm);
            accm = fmaf(a.dsq[2 * p + 1], byte2f(mn4, 2 * p + 1), accm);
        }
        return dm.x * accd - dm.y * accm;
    } else if constexpr (TYPE == T_Q5_K) {         // vecdotq.cuh:415-444: as Q4_K with bit (sub-block) of qh[l] as the code's fifth bit
        const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&w.hd[0]));
        const uint32_t h1 = (uint32_t)w.hd[1], h2 = (uint32_t)w.hd[2], h3 = (uint32_t)w.hd[3];
        const uint32_t sc4 = hh ? ((h3 & 0x0F0F0F0Fu) | ((h1 >> 2) & 0x30303030u)) : (h1 & 0x3F3F3F3Fu);
        const uint32_t mn4 = hh ? (((h3 >> 4) & 0x0F0F0F0Fu) | ((h2 >> 2) & 0x30303030u)) : (h2 & 0x3F3F3F3Fu);
        float accd = 0.0f, accm = 0.0f;
#pragma unroll
        for (int p = 0; p < 2; ++p) {
            int s0 = 0, s1 = 0;
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const uint32_t q = (uint32_t)w.w[8 * p + j];
                const uint32_t hb = hh ? ((uint32_t)w.qh[j] >> 4) : (uint32_t)w.qh[j];        // this half's four sub-block bits at bits 0..3
                const uint32_t b0 = (hb << (4 - 2 * p)) & 0x10101010u, b1 = (hb << (3 - 2 * p)) & 0x10101010u;
                s0 = dp4a_us((q & 0x0F0F0F0Fu) | b0, a.q[2 * p][j], s0);
                s1 = dp4a_us(((q >> 4) & 0x0F0F0F0Fu) | b1, a.q[2 * p + 1][j], s1);
            }
            if constexpr (DBG) { dbg[2 * p] = s0; dbg[2 * p + 1] = s1; }
            accd = fmaf(a.d8[2 * p], i2f_small(s0) * byte2f(sc4, 2 * p), accd);
            accd = fmaf(a.d8[2 * p + 1], i2f_small(s1) * byte2f(sc4, 2 * p + 1), accd);
            accm = fmaf(a.dsq[2 * p], byte2f(mn4, 2 * p), accm);
            accm = fmaf(a.dsq[2 * p + 1], byte2f(mn4, 2 * p + 1), accm);
        }
        return dm.x * accd - dm.y * accm;
    } else if constexpr (TYPE == T_Q4_0) {         // vecdotq.cuh:25-44: d_w * d8 * (sum nibble q_x - 8 sum q_x) per 32-block
        float acc = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int s0 = 0, s1 = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t q = (uint32_t)word_at_half<18>(w.w, 9 * j + 1 + 2 * k);
                s0 = dp4a_us(q & 0x0F0F0F0Fu, a.q[j][k], s0);
                s1 = dp4a_us(q & 0xF0F0F0F0u, a.q[j][4 + k], s1);          // 16 x nibble, divided out below
            }
            const int s = s0 + (s1 >> 4);
            if constexpr (DBG) dbg[j] = s;
            acc = fmaf(half_bits_to_float(halfword<18>(w.w, 9 * j)) * a.d8[j], i2f_small(s + ((a.m32a[j] + a.m32b[j]) >> 2)), acc);
        }
        return acc;
    } else if constexpr (TYPE == T_Q4_1) {         // vecdotq.cuh:49-80 (the block minimum times d8 * sum q from the layout's tail)
        float acc = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&w.w[5 * j]));
            int s0 = 0, s1 = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t q = (uint32_t)w.w[5 * j + 1 + k];
                s0 = dp4a_us(q & 0x0F0F0F0Fu, a.q[j][k], s0);
                s1 = dp4a_us(q & 0xF0F0F0F0u, a.q[j][4 + k], s1);
            }
            const int s = s0 + (s1 >> 4);
            if constexpr (DBG) dbg[j] = s;
            acc = fmaf(dm.x * a.d8[j], i2f_small(s), fmaf(dm.y, a.dsq[j], acc));
        }
        return acc;
    } else if constexpr (TYPE == T_Q5_0 || TYPE == T_Q5_1) {      // vecdotq.cuh:82-150
        float acc = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            float dw, mw = 0.0f; uint32_t qh, qs[4];
            if constexpr (TYPE == T_Q5_0) {
                dw = half_bits_to_float(halfword<22>(w.w, 11 * j));
                qh = (uint32_t)word_at_half<22>(w.w, 11 * j + 1);
#pragma unroll
                for (int k = 0; k < 4; ++k) qs[k] = (uint32_t)word_at_half<22>(w.w, 11 * j + 3 + 2 * k);
            } else {
                const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&w.w[6 * j]));
                dw = dm.x; mw = dm.y;
                qh = (uint32_t)w.w[6 * j + 1];
#pragma unroll
                for (int k = 0; k < 4; ++k) qs[k] = (uint32_t)w.w[6 * j + 2 + k];
            }
            int s = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t b0 = ((((qh >> (4 * k)) & 0xFu) * 0x00204081u) & 0x01010101u) << 4;
                const uint32_t b1 = ((((qh >> (16 + 4 * k)) & 0xFu) * 0x00204081u) & 0x01010101u) << 4;
                s = dp4a_us((qs[k] & 0x0F0F0F0Fu) | b0, a.q[j][k], s);
                s = dp4a_us(((qs[k] >> 4) & 0x0F0F0F0Fu) | b1, a.q[j][4 + k], s);
            }
            if constexpr (DBG) dbg[j] = s;
            if constexpr (TYPE == T_Q5_0) acc = fmaf(dw * a.d8[j], i2f_small(s + ((a.m32a[j] + a.m32b[j]) >> 1)), acc);
            else acc = fmaf(dw * a.d8[j], i2f_small(s), fmaf(mw, a.dsq[j], acc));
        }
        return acc;
    } else if constexpr (TYPE == T_IQ4_NL) {
        float acc = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int s = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int vlo, vhi; iq4_lookup8((uint32_t)word_at_half<18>(w.w, 9 * j + 1 + 2 * k), vlo, vhi);
                s = dp4a_ss(vlo, a.q[j][k], s); s = dp4a_ss(vhi, a.q[j][4 + k], s);
            }
            if constexpr (DBG) dbg[j] = s;
            acc = fmaf(half_bits_to_float(halfword<18>(w.w, 9 * j)) * a.d8[j], i2f_small(s), acc);
        }
        return acc;
    } else if constexpr (TYPE == T_IQ4_XS) {
        const float d = half_bits_to_float((uint32_t)w.hd[0] & 0xFFFFu);
        const uint32_t sh = (uint32_t)w.hd[0] >> 16, sl = (uint32_t)w.hd[1];
        float accx = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int ib = 4 * hh + j;
            const int ls = (int)(((sl >> (4 * ib)) & 0xFu) | (((sh >> (2 * ib)) & 3u) << 4)) - 32;
            int s = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                int vlo, vhi; iq4_lookup8((uint32_t)w.q[4 * j + k], vlo, vhi);
                s = dp4a_ss(vlo, a.q[j][k], s); s = dp4a_ss(vhi, a.q[j][4 + k], s);
            }
            if constexpr (DBG) dbg[j] = s;
            accx = fmaf(a.d8[j] * (float)ls, i2f_small(s), accx);
        }
        return d * accx;
    } else if constexpr (TYPE == T_Q2_K) {         // vecdotq.cuh:218-243: this half's blocks j = 0..3 use bit pair j of the 32 code bytes
        const float2 dm = __half22float2(*reinterpret_cast<const __half2 *>(&w.dm));
        float accd = 0.0f, accm = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t scw = (uint32_t)w.sc[j >> 1] >> (16 * (j & 1));
            const int sa = scw & 0xF, ma = (scw >> 4) & 0xF, sb = (scw >> 8) & 0xF, mb = (scw >> 12) & 0xF;
            int s0 = 0, s1 = 0;
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                s0 = dp4a_us(((uint32_t)w.q[k] >> (2 * j)) & 0x03030303u, a.q[j][k], s0);
                s1 = dp4a_us(((uint32_t)w.q[4 + k] >> (2 * j)) & 0x03030303u, a.q[j][4 + k], s1);
            }
            if constexpr (DBG) dbg[j] = s0 + s1;
            accd = fmaf(a.d8[j], i2f_small(s0 * sa + s1 * sb), accd);
            accm = fmaf(a.d8[j], i2f_small(-(a.m32a[j] >> 5) * ma - (a.m32b[j] >> 5) * mb), accm);
        }
        return dm.x * accd - dm.y * accm;
    } else if constexpr (TYPE == T_Q3_K) {         // vecdotq.cuh:301-331
        const float d = half_bits_to_float(w.d);
        float acc3 = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int scv[2];
#pragma unroll
            for (int h2 = 0; h2 < 2; ++h2) {       // scale 8hh + 2j + h2: low nibble word (2j + h2) / 4, byte (2j + h2) % 4, second half: high nibbles
                const int il = 2 * j + h2;
                const uint32_t lo = ((uint32_t)w.sc[il >> 2] >> (8 * (il & 3) + 4 * hh)) & 0xFu;
                const uint32_t hi = ((uint32_t)w.sc[2] >> (8 * (il & 3) + 2 * (2 * hh + (il >> 2)))) & 3u;
                scv[h2] = (int)(lo | (hi << 4)) - 32;
            }
            int s0 = a.m32a[j] >> 3, s1 = a.m32b[j] >> 3;      // -4 * sum q8 of each half
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t v0 = (((uint32_t)w.q[k] >> (2 * j)) & 0x03030303u) | ((((uint32_t)w.hm[k] >> (4 * hh + j)) & 0x01010101u) << 2);
                const uint32_t v1 = (((uint32_t)w.q[4 + k] >> (2 * j)) & 0x03030303u) | ((((uint32_t)w.hm[4 + k] >> (4 * hh + j)) & 0x01010101u) << 2);
                s0 = dp4a_us(v0, a.q[j][k], s0);
                s1 = dp4a_us(v1, a.q[j][4 + k], s1);
            }
            if constexpr (DBG) dbg[j] = s0 + s1;
            acc3 = fmaf(a.d8[j], i2f_small(s0 * scv[0] + s1 * scv[1]), acc3);
        }
        return d * acc3;
    } else if constexpr (TYPE == T_Q8_0) {         // vecdotq.cuh:153-165: d_w * d8 * sum q_w q_x per 32-block
        float acc = 0.0f;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            int s = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) s = dp4a_ss(word_at_half<34>(w.w, 17 * j + 1 + 2 * k), a.q[j][k], s);
            if constexpr (DBG) dbg[j] = s;
            acc = fmaf(half_bits_to_float(halfword<34>(w.w, 17 * j)) * a.d8[j], i2f_small(s), acc);
        }
        return acc;
    } else {
        const float d = half_bits_to_float(w.d);
        float acc6 = 0.0f;
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            int sa = a.m32a[g], sb = a.m32b[g];        // -32 * sum q: folds the -32 offset of the 6-bit code
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                const uint32_t l4 = (uint32_t)w.ql[8 * (g & 1) + j], h4 = (uint32_t)w.qh[j];
                const uint32_t nib = (g < 2 ? l4 : (l4 >> 4)) & 0x0F0F0F0Fu;
                const uint32_t top = (g == 0 ? (h4 << 4) : g == 1 ? (h4 << 2) : g == 2 ? h4 : (h4 >> 2)) & 0x30303030u;
                if (j < 4) sa = dp4a_us(nib | top, a.q[g][j], sa); else sb = dp4a_us(nib | top, a.q[g][j], sb);
            }
            const int sca = (int)(int8_t)((uint32_t)w.sc[g >> 1] >> (16 * (g & 1)));
            const int scb = (int)(int8_t)((uint32_t)w.sc[g >> 1] >> (16 * (g & 1) + 8));
            if constexpr (DBG) dbg[g] = sa + sb;
            acc6 = fmaf(a.d8[g], i2f_small(sa * sca + sb * scb), acc6);
        }
        return d * acc6;
    }
}

template <int TYPE, int U, bool DBG, bool NORM>
__global__ void __launch_bounds__(288, 2) mmvqr_kernel(const mmvqr_params p) {
    constexpr int IB = item_traits<TYPE>::BYTES;
    extern __shared__ __align__(128) uint8_t smem_raw[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int NW = p.TPR > 32 ? p.TPR >> 5 : 1;            // partials per row
    const int RPP = 256 >> p.lgTPR;                        // rows per pass

    uint8_t * stages = smem_raw;
    uint8_t * act = smem_raw + (size_t)p.nstages * p.stage_bytes;                                  // BPR items of 400 bytes
    float * xs = reinterpret_cast<float *>(act + (size_t)p.BPR * ACT_ITEM);                        // NORM: the K floats to normalise
    float * ws = xs + (NORM ? p.K : 0);                                                            // NORM: the norm weights
    float * partials = ws + (NORM ? p.K : 0);                                                      // [nstages][R][NW]
    const int part_stage = p.R * NW;
    uint64_t * bars = reinterpret_cast<uint64_t *>(partials + (((size_t)p.nstages * part_stage + 1) & ~(size_t)1));
    const uint32_t full0 = smem_u32(bars), empty0 = full0 + 8 * p.nstages, actbar = empty0 + 8 * p.nstages, wbar = actbar + 8;
    float * npart = reinterpret_cast<float *>(bars + 2 * p.nstages + 2);     // 32 floats: sums of squares per 32 float4s
    float * rsd = npart + 32;                                                // resid_rows floats: this CTA's residual values
    float * csn = rsd + p.resid_rows;                                        // rope_rows floats: (cos, sin) of this CTA's row pairs

    int cta = blockIdx.x, g = 0;
    while (g + 1 < p.ngroup && cta >= p.cta0[g + 1]) ++g;
    cta -= p.cta0[g];
    const uint8_t * Wc = p.Wg[g];
    float * Dc = p.Dg[g];
    const bool pair_ids = p.sm_act != nullptr && p.ids != nullptr;
    const uint8_t * Wc1 = pair_ids ? p.Wup : p.Wg[1];                         // pair mode: the second matrix of this CTA
    float * Dc1 = pair_ids ? p.Du[g] : p.Dg[1];
    float * act_o = p.sm_act ? p.sm_act + (pair_ids ? g * p.sm_slot : 0) : nullptr, * out_o = p.sm_out ? p.sm_out + (pair_ids ? g * p.sm_slot : 0) : nullptr;
    uint8_t * xq_o = p.sm_xq ? p.sm_xq + (pair_ids ? g * p.sm_xq_col : 0) : nullptr;
    const bool pair = p.sm_act != nullptr;     // gate / up pair mode: this CTA's rows of matrix 0, then the same rows of matrix 1
    const int unit = pair ? 32 : p.rope_on ? 2 : 1;        // rope: rows are dealt in pairs, pair mode: in 32-row blocks (rows_qg / rows_remg count units)
    const int row_lo = unit * (cta * p.rows_qg[g] + min(cta, p.rows_remg[g]));
    const int nrows = unit * (p.rows_qg[g] + (cta < p.rows_remg[g] ? 1 : 0));
    // Stage s of this CTA covers rows [srow(s), srow(s) + snr(s)) of its range.  The first three stages are short (r0, r1,
    // r2 rows): bulk copies in flight share the SM's bandwidth, so equal stages all land together after the whole batch
    // has streamed (the first dot product of a 66 MB gate/up GEMV started 5-8 us after its loads were issued,
    // profiles/r01_ktrace_tg_v5.txt); with a ramp the first rows arrive after ~1 us and the multiply overlaps the rest.
    const int rampC = p.r0 + p.r1 + p.r2;
    auto srow = [&](int s) { return s == 0 ? 0 : s == 1 ? p.r0 : s == 2 ? p.r0 + p.r1 : rampC + (s - 3) * p.R; };
    auto snr = [&](int s) { return min(s == 0 ? p.r0 : s == 1 ? p.r1 : s == 2 ? p.r2 : p.R, nrows - srow(s)); };
    const int nst = nrows <= p.r0 ? 1 : nrows <= p.r0 + p.r1 ? 2 : nrows <= rampC ? 3 : 3 + (nrows - rampC + p.R - 1) / p.R;
    const int nsv = pair ? 2 * nst : nst;      // stages this CTA streams (pair mode: sv < nst matrix 0, else matrix 1)

    if (tid == 0) {
        for (int s = 0; s < p.nstages; ++s) { mbar_init(full0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 8); }
        mbar_init(actbar, 1);
        mbar_init(wbar, 1);
        mbar_fence_init();
    }
    __syncthreads();
    pdl_trigger();            // the next kernel of the step may become resident and start its own weight stream

    if (warp == 0) {
        const uint64_t pol = policy_evict_first();
        auto issue = [&](int sv) {   // lane 0 only
            const int slot = sv % p.nstages;
            const int m = pair && sv >= nst ? 1 : 0, s = sv - m * nst;
            const int r0 = row_lo + srow(s), nr = snr(s);
            const uint32_t bytes = (uint32_t)nr * p.rb, bar = full0 + 8 * slot;
            mbar_expect_tx(bar, bytes);
            bulk_g2s_stream(smem_u32(stages + (size_t)slot * p.stage_bytes), (m ? Wc1 : Wc) + (int64_t)r0 * p.rb, bytes, bar, pol);
        };
        // NORM kernels follow another mat-vec directly (no early residency), so their row must not queue behind the whole
        // weight prefetch in the SM's copy engine: norm weights and the first stage before the wait, the row right after
        // it, the other stages behind the row.  (296 CTAs reading the same 16 KB with ordinary loads took 2-8 us, the bulk
        // copies of the same lines about 1 us: profiles/r01_ktrace_tg_v6.txt.)
        if (lane == 0 && !p.ids) {
            if (NORM) { mbar_expect_tx(wbar, (uint32_t)p.K * 4); bulk_g2s(smem_u32(ws), p.nw, (uint32_t)p.K * 4, wbar); }
            for (int s = 0; s < min(nsv, (NORM || p.stagger) ? 1 : p.nstages); ++s) issue(s);
            if (!NORM && p.stagger && nsv > 1) { mbar_wait(full0, 0); for (int s = 1; s < min(nsv, p.nstages); ++s) issue(s); }
        }
        const int rmode = p.rope_on ? p.rmode[g] : 0;
        if (rmode == 1 || rmode == 2) {
            // cos / sin of this CTA's row pairs: depends on the position only (a graph input), so it is computed while the
            // previous kernel is still running and the weights are in flight
            const int hd = p.rhd[g], posv = p.rpos[0];
            for (int o2 = lane; 2 * o2 < nrows; o2 += 32) {
                const int d = (row_lo + 2 * o2) % hd;
                float c = 1.0f, sn = 0.0f;
                if (d < p.rp.n_dims) rope_cs(p.rp, d >> 1, posv, p.rff, c, sn);
                csn[2 * o2] = c; csn[2 * o2 + 1] = sn;
            }
            __syncwarp();
        }
        pdl_wait();           // the activations come from the previous kernel; dst may still be read by it
        if (lane == 0) hs_wait_acts(p.hs);      // row split, peer side: the main device's broadcast has landed
        if (p.ids) { Wc += (int64_t)p.ids[g] * p.w_expert; Wc1 += (int64_t)p.ids[g] * p.w_expert; }      // the expert the router chose for slot g (known only now)
        if (lane == 0) {      // one coalesced copy of the column; the consumers pick their registers out of shared memory
            const uint32_t ab = NORM ? (uint32_t)p.K * 4 : (uint32_t)p.BPR * ACT_ITEM;
            mbar_expect_tx(actbar, ab);
            bulk_g2s(NORM ? smem_u32(xs) : smem_u32(act), NORM ? (const void *)p.nx : (const void *)(p.Xq + (size_t)(p.ids ? p.xq_col[g] : 0) * ab), ab, actbar);
            if (NORM) for (int s = 1; s < min(nsv, p.nstages); ++s) issue(s);
            if (p.ids) for (int s = 0; s < min(nsv, p.nstages); ++s) issue(s);
        }
        if (p.resid) {        // this CTA's rows of the residual, fetched while the first stage is being multiplied
            for (int o = lane; o < nrows; o += 32) rsd[o] = p.resid[row_lo + o];
            __syncwarp();
        }
        for (int sv = 0; sv < nsv; ++sv) {
            const int slot = sv % p.nstages;
            mbar_wait(empty0 + 8 * slot, (sv / p.nstages) & 1);
            const int mtx = pair && sv >= nst ? 1 : 0, s = sv - mtx * nst;
            const int r0 = row_lo + srow(s), nr = snr(s);
            const float * ps = partials + (size_t)slot * part_stage;
            if (pair) {
                // rsd: this CTA's gate sums, csn: its products (both nrows floats of shared memory)
                for (int o = lane; o < nr; o += 32) {
                    float sum = 0.0f;
                    for (int wv = 0; wv < NW; ++wv) sum += ps[o * NW + wv];
                    const int n = r0 + o, l = n - row_lo;
                    if (mtx == 0) { Dc[n] = sum; rsd[l] = sum; }
                    else {
                        const float av = silu_fn(rsd[l]), y = av * sum;
                        Dc1[n] = sum;                  // act / out may sit on the gate / up results: these stores come last, act before out
                        act_o[n] = av; out_o[n] = y;
                        csn[l] = y;
                    }
                }
            } else if (rmode == 0) {
                for (int o = lane; o < nr; o += 32) {
                    float sum = 0.0f;
                    for (int wv = 0; wv < NW; ++wv) sum += ps[o * NW + wv];
                    Dc[r0 + o] = sum;
                    if (p.resid) p.dsum[r0 + o] = sum + rsd[r0 - row_lo + o];
                }
            } else {
                // q / k: rotate the adjacent pair (2i, 2i+1) -- the partner's sum sits in the neighbouring lane (r0 and nr are
                // even) -- and store the result (f32 tensor of the ROPE node, K row of the cache); v: the row into the cache
                char * cell = p.rslot[g] ? *p.rslot[g] : p.rcell[g];
                for (int o0 = 0; o0 < nr; o0 += 32) {
                    const int o = o0 + lane, n = r0 + o;
                    const bool valid = o < nr;
                    float sum = 0.0f;
                    if (valid) for (int wv = 0; wv < NW; ++wv) sum += ps[o * NW + wv];
                    if (rmode == 3) {
                        if (valid) { Dc[n] = sum; store_as(cell + (size_t)n * p.rcesz[g], p.rctype[g], sum); }
                        continue;
                    }
                    const float other = __shfl_xor_sync(0xffffffffu, sum, 1);
                    if (valid) {
                        float y = sum;
                        if ((n % p.rhd[g]) < p.rp.n_dims) {
                            const int lo = (n - row_lo) & ~1;
                            float ya, yb;
                            rope_pair((n & 1) ? other : sum, (n & 1) ? sum : other, csn[lo], csn[lo + 1], ya, yb);
                            y = (n & 1) ? yb : ya;
                        }
                        p.rout[g][n] = y;
                        if (rmode == 2) store_as(cell + (size_t)n * p.rcesz[g], p.rctype[g], y);
                    }
                }
            }
            __syncwarp();
            if (lane == 0 && sv + p.nstages < nsv) issue(sv + p.nstages);
            if (sv == 0) ktrace_event(2);       // first stage consumed and stored
        }
        ktrace_event(3);                       // all rows of CTA 0 stored
        if (pair && xq_o) {
            // the q8 activations of this CTA's 32-row blocks, four blocks per pass (an 8-lane group each)
            __syncwarp();
            for (int b0 = 0; 32 * b0 < nrows; b0 += 4) {
                const int bl = b0 + (lane >> 3);
                const bool in = 32 * bl < nrows;
                float4 y = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                if (in) { const float * yp = csn + 32 * bl + 4 * (lane & 7); y = make_float4(yp[0], yp[1], yp[2], yp[3]); }
                quant_block_store(y, in ? (row_lo >> 5) + bl : 0, lane & 7, in, xq_o);
            }
        }
        if (p.hs.counter) {                    // row split: the last CTA of the launch lowers / raises / awaits the hand-over flags
            __syncwarp();
            if (lane == 0) hs_finish(p.hs, gridDim.x);
        }
    } else {
        const int ct = tid - 32;
        const int h = ct & (p.TPR - 1), rsub = ct >> p.lgTPR;
        const bool active = h < 2 * p.BPR;
        const int b = h >> 1, hh = h & 1;
        const int wir = h >> 5;                              // warp inside the row (TPR > 32)
        const int red = min(p.TPR, 32);
        hacts a;
        if constexpr (NORM) {
            // RMS_NORM + MUL + q8 quantisation of the column, the arithmetic of add_rms_norm_mul_kernel<1> (one float4 per
            // thread of a 1024-thread CTA there; here thread ct takes float4s ct, ct + 256, .. so that a warp still holds
            // 32 consecutive float4s and the butterfly sums, the order of the partials and the 8-lane quantiser groups
            // are the same)
            const int n4 = p.K >> 2;
            const int cw = warp - 1;
            // Rolled loops on purpose: this code runs once per launch out of a cold instruction cache while the weight copies
            // saturate L2, so its cost is its SIZE (the unrolled form, 1600 instructions between the two barriers, took 7 us;
            // see profiles/r02_norm_prologue.md); the row and the weights are re-read from shared memory instead of kept in registers.
            mbar_wait(wbar, 0);                     // norm weights: copied before the PDL wait
            mbar_wait(actbar, 0);                   // the row is in shared memory (copied after the producer's PDL wait)
            ktrace_event(4, 32);
#pragma unroll 1
            for (int k = 0; k < 4; ++k) {
                const int i = ct + 256 * k;
                if (i < n4) {
                    const float sq = warp_sum(sumsq4(reinterpret_cast<const float4 *>(xs)[i]));
                    if (lane == 0) npart[8 * k + cw] = sq;
                }
            }
            ktrace_event(5, 32);
            asm volatile("bar.sync 1, 256;" ::: "memory");
            float tot = 0.0f;
            if ((n4 >> 5) == 32) {                                     // K = 4096: loads batched, the same sequential order of additions
#pragma unroll
                for (int w2 = 0; w2 < 16; ++w2) { const float2 t2 = reinterpret_cast<const float2 *>(npart)[w2]; tot += t2.x; tot += t2.y; }
            } else {
                for (int wi = 0; wi < (n4 >> 5); ++wi) tot += npart[wi];
            }
            const float scale = rms_scale(tot, p.K, p.neps);
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
                        const float4 w0 = reinterpret_cast<const float4 *>(ws)[i], w1 = reinterpret_cast<const float4 *>(ws)[i + 1];
                        const float4 n0 = make_float4(v0.x * scale, v0.y * scale, v0.z * scale, v0.w * scale);
                        const float4 n1 = make_float4(v1.x * scale, v1.y * scale, v1.z * scale, v1.w * scale);
                        y0[k] = make_float4(n0.x * w0.x, n0.y * w0.y, n0.z * w0.z, n0.w * w0.w);
                        y1[k] = make_float4(n1.x * w1.x, n1.y * w1.y, n1.z * w1.z, n1.w * w1.w);
                        if (writer) {
                            reinterpret_cast<float4 *>(p.nnorm)[i] = n0; reinterpret_cast<float4 *>(p.nnorm)[i + 1] = n1;
                            reinterpret_cast<float4 *>(p.nout)[i] = y0[k]; reinterpret_cast<float4 *>(p.nout)[i + 1] = y1[k];
                        }
                    }
                }
                quant_block_store8<2>(y0, y1, jb, ct & 3, in, act, writer ? p.nxq : nullptr);
            }
            asm volatile("bar.sync 1, 256;" ::: "memory");
            ktrace_event(6, 32);
        } else {
            mbar_wait(actbar, 0);
            ktrace_event(6, 32);
        }
        if (active) load_hacts(smem_u32(act) + (uint32_t)b * ACT_ITEM + hh * (4 * ACT_BLK), a);
        const uint32_t ioff = (uint32_t)b * IB;
        const int nb32 = p.K >> 5;
        for (int sv = 0; sv < nsv; ++sv) {
            const int slot = sv % p.nstages;
            const int s = pair && sv >= nst ? sv - nst : sv;
            const int r0 = row_lo + srow(s), nr = snr(s);
            mbar_wait(full0 + 8 * slot, (sv / p.nstages) & 1);
            if (sv == 0) ktrace_event(7, 32);       // first weights have landed
            const uint32_t sbase = smem_u32(stages + (size_t)slot * p.stage_bytes) + ioff;
            float * pw = partials + (size_t)slot * part_stage;
            // U passes (rows rr, rr + RPP, ..) in flight per thread: independent shared loads and dp4a chains
            for (int rr = rsub; rr - rsub < nr; rr += U * RPP) {
                hweights<TYPE> w[U];
                float x[U];
#pragma unroll
                for (int u = 0; u < U; ++u)
                    if (active && rr + u * RPP < nr) load_hweights<TYPE>(sbase + (uint32_t)(rr + u * RPP) * p.rb, hh, w[u]);
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    x[u] = 0.0f;
                    if (active && rr + u * RPP < nr)
                        x[u] = half_dot<TYPE, DBG>(w[u], a, hh, DBG ? p.dbg_sumi + (int64_t)(r0 + rr + u * RPP) * nb32 + 8 * b + 4 * hh : nullptr);
                }
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    if (o < red) {
#pragma unroll
                        for (int u = 0; u < U; ++u) x[u] += __shfl_xor_sync(0xffffffffu, x[u], o);
                    }
                }
                if ((lane & (red - 1)) == 0) {
#pragma unroll
                    for (int u = 0; u < U; ++u)
                        if (rr + u * RPP < nr) pw[(rr + u * RPP) * NW + wir] = x[u];
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(empty0 + 8 * slot);
        }
    }
}

constexpr int MMVQR_SKIP = -777;
struct mmvqr_launch_cfg { dim3 grid; size_t smem; int pdl; int U; };
template <int TYPE> int mmvqr_launch_type(const mmvqr_params & p, const mmvqr_launch_cfg & c, cudaStream_t stream);

template <int TYPE, int U, bool DBG, bool NORM>
static int mmvqr_launch_one(const mmvqr_params & p, const mmvqr_launch_cfg & c, cudaStream_t stream) {
    auto kern = mmvqr_kernel<TYPE, U, DBG, NORM>;
    static int smem_set[16] = {0};   // per device
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 16 && !smem_set[dev]) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 232448));
        smem_set[dev] = 1;
    }
    cudaLaunchConfig_t lc = {};
    lc.gridDim = c.grid; lc.blockDim = dim3(288, 1, 1); lc.dynamicSmemBytes = c.smem; lc.stream = stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = c.pdl ? 1 : 0;
    B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, kern, p));
    return 0;
}

#define B200_MMVQR_DEFINE(TYPE_) \
    template <> int mmvqr_launch_type<TYPE_>(const mmvqr_params & p, const mmvqr_launch_cfg & c, cudaStream_t stream) { \
        if (p.nx) return mmvqr_launch_one<TYPE_, 1, false, true>(p, c, stream); \
        if (p.dbg_sumi) return c.U == 2 ? mmvqr_launch_one<TYPE_, 2, true, false>(p, c, stream) : mmvqr_launch_one<TYPE_, 1, true, false>(p, c, stream); \
        return c.U == 2 ? mmvqr_launch_one<TYPE_, 2, false, false>(p, c, stream) : mmvqr_launch_one<TYPE_, 1, false, false>(p, c, stream); \
    }

}  // namespace b200
