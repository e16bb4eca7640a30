// fattn_mma.cuh -- decode attention over a long f16 KV cache with warp-level tensor-core tiles (included by fattn.cu).
//
// Same operator and split-KV structure as fattn_vec_gqa_kernel (reference: ggml-cuda/fattn-common.cuh:613-672 split + combine,
// fattn-mma-f16.cuh for the idea of feeding a handful of query rows to mma tiles): all queries that share a KV head (GQA heads x
// rows, up to 16) form the M = 16 side of m16n8k16 tiles, so a K row and a V row are read once for the whole group.
//
// Why: the vector kernel spends ~2900 warp instructions per 16 KB of cache (one FMA per query, key element and lane) and was
// instruction bound at 1.1 of 6.5 TB/s at 32k context (profiles/r02_fattn_gqa_32k_ncu.md).  Here 32 positions (16 KB of K + V at
// D = 128) cost 64 mma + 32 ldmatrix + the softmax of a 16 x 32 tile: ~250 instructions.
//
//   CTA = 4 warps, one per SM (192 KB of shared memory: 3 stages of 16 KB per warp).  Each warp runs its own cp.async pipeline over
//   its chunks of 32 positions (chunk c of the split goes to warp c % 4): no CTA-wide barrier in the main loop.  Rows of a chunk sit
//   in shared memory with the 16-byte pieces xor-swizzled by (row & 7): ldmatrix reads them conflict free.
//   S = Q K^T: A = Q (f16, 8 k-steps held in registers for the whole kernel), B = K rows as stored (ldmatrix).
//   online softmax on the accumulator fragments (thread = 2 rows x 8 positions), P packed to f16 in registers as the A operand of
//   O += P V with B = V rows through ldmatrix.trans.  Positions past n_kv are zero-filled by cp.async (src-size 0) and masked.
//   The four warps' (M, L, O) are merged through shared memory; splits through fattn_combine_kernel as before.
#pragma once

namespace b200 {

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t (&r)[4]) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t (&r)[4]) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma_f16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_h2(float x, float y) {
    const __half2 h = __floats2half2_rn(x, y);
    return *reinterpret_cast<const uint32_t *>(&h);
}
__device__ __forceinline__ void cp_async16_zfill(uint32_t dst, const void * src, uint32_t src_bytes) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" :: "r"(dst), "l"(src), "r"(src_bytes) : "memory");
}

constexpr int FMMA_NW = 4, FMMA_NS = 3, FMMA_CH = 32;
constexpr int FMMA_MROW = 80;                              // bytes per mask row of a stage: 32 halves + padding (conflict-free 4-byte reads)
// Q8 (q8_0 cache): the stages hold the raw rows (D / 32 blocks of {half d; int8 q[32]}); each warp expands the chunk it is about to
// multiply into one f16 tile pair of its own (shared to shared, d * q rounded to f16: what the reference's tensor-core attention
// does to the whole cache up front, fattn-common.cuh:729-757), then runs the f16 path on it.
template <int D, bool Q8> constexpr size_t fmma_smem_bytes() {
    return 16 * 2 * D + (size_t)FMMA_NW * (FMMA_NS * (2 * FMMA_CH * (Q8 ? D / 32 * 34 : 2 * D) + 16 * FMMA_MROW) + (Q8 ? 2 * FMMA_CH * 2 * D : 0));
}
__device__ __forceinline__ void sts128_u4(uint32_t addr, uint4 v) {
    asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" :: "r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
template <int BYTES> __device__ __forceinline__ void cp_async_small_zfill(uint32_t dst, const void * src, uint32_t src_bytes) {
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2, %3;" :: "r"(dst), "l"(src), "n"(BYTES), "r"(src_bytes) : "memory");
}
// two int8 (low 16 bits of w) times the block scale -> f16x2: 0x6400 | (q ^ 0x80) is the f16 1024 + 128 + q, exactly
__device__ __forceinline__ uint32_t q8_pair_f16(uint32_t w, __half2 d2) {
    const uint32_t u = __byte_perm(w ^ 0x8080u, 0x64646464u, 0x4140);
    const __half2 x = __hsub2(*reinterpret_cast<const __half2 *>(&u), __half2(__ushort_as_half((unsigned short)0x6480), __ushort_as_half((unsigned short)0x6480)));
    const __half2 y = __hmul2(x, d2);
    return *reinterpret_cast<const uint32_t *>(&y);
}

template <int D, bool Q8>
__global__ void __launch_bounds__(32 * FMMA_NW, 1) fattn_mma_gqa_kernel(const fa_params p) {
    pdl_trigger(); pdl_wait();
    constexpr int NW = FMMA_NW, NS = FMMA_NS, CH = FMMA_CH;
    constexpr int ROWB = 2 * D, PPR = ROWB / 16;           // bytes and 16-byte pieces per K / V row
    constexpr int CHB = CH * ROWB;                         // an f16 K (or V) chunk
    constexpr int RAWB = Q8 ? D / 32 * 34 : ROWB, RCH = CH * RAWB;      // a row / a chunk as staged (raw q8_0 blocks, or the f16 rows themselves)
    constexpr int STB = 2 * RCH + 16 * FMMA_MROW;          // a stage = K chunk + V chunk + 16 mask rows
    constexpr int PB = D == 128 ? 8 : 4, RPR = RAWB / PB;  // q8_0: copy granule (the rows are only 8- / 4-byte aligned) and granules per row
    constexpr int KS = D / 16, NB = D / 8;                 // k-steps of Q K^T, n-blocks of O
    extern __shared__ __align__(128) uint8_t fm_smem[];
    __shared__ float wm[NW][16], wl[NW][16];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int r = lane >> 2, t = lane & 3;                 // accumulator fragment: rows r, r + 8; columns 2t, 2t + 1 of every n-block
    const uint32_t qs = smem_u32(fm_smem);
    const uint32_t wb = qs + 16 * ROWB + (uint32_t)warp * NS * STB;
    const uint32_t ft = qs + 16 * ROWB + (uint32_t)NW * NS * STB + (uint32_t)warp * 2 * CHB;      // Q8: this warp's f16 tiles (K, V)

    const int ikv = blockIdx.y % (p.n_head / p.rk2), iq3 = blockIdx.y / (p.n_head / p.rk2), split = blockIdx.z;
    const int ngq = p.N * p.rk2;                           // queries that share this KV head
    const int g0 = blockIdx.x * 16;
    // Q -> f16, swizzled rows in shared memory (rows past the group are zero)
    for (int i = tid; i < 16 * PPR; i += 32 * NW) {
        const int g = i / PPR, c = i % PPR, gq = g0 + g;
        uint4 u = make_uint4(0u, 0u, 0u, 0u);
        if (gq < ngq) {
            const float * qp = reinterpret_cast<const float *>(p.q + (int64_t)(gq / p.rk2) * p.nbq1 + (int64_t)(ikv * p.rk2 + gq % p.rk2) * p.nbq2 + (int64_t)iq3 * p.nbq3) + 8 * c;
            const float4 x0 = *reinterpret_cast<const float4 *>(qp), x1 = *reinterpret_cast<const float4 *>(qp + 4);
            u = make_uint4(pack_h2(x0.x, x0.y), pack_h2(x0.z, x0.w), pack_h2(x1.x, x1.y), pack_h2(x1.z, x1.w));
        }
        *reinterpret_cast<uint4 *>(fm_smem + g * ROWB + ((c ^ (g & 7)) << 4)) = u;
    }
    const char * kbase = p.k + (int64_t)ikv * p.nbk2 + (int64_t)(iq3 / p.rk3) * p.nbk3;
    const char * vbase = p.v + (int64_t)ikv * p.nbv2 + (int64_t)(iq3 / p.rv3) * p.nbv3;
    const int ntiles = (p.n_kv + CH - 1) / CH;
    const int t0 = split * p.tiles_per_split, t1 = min(ntiles, t0 + p.tiles_per_split);
    const int nch = t1 > t0 + warp ? (t1 - t0 - warp + NW - 1) / NW : 0;      // chunks of this warp: t0 + warp + NW * c

    auto issue = [&](int c) {                              // all lanes: 2 * CH * PPR / 32 copies of 16 bytes
        const int pos0 = CH * (t0 + warp + NW * c);
        const uint32_t sb = wb + (uint32_t)(c % NS) * STB;
        if constexpr (Q8) {
#pragma unroll
            for (int i = 0; i < CH * RPR / 32; ++i) {
                const int q = 32 * i + lane, row = q / RPR, pc = q % RPR;
                const bool in = pos0 + row < p.n_kv;
                const int64_t pr = in ? pos0 + row : 0;
                const uint32_t off = (uint32_t)row * RAWB + (uint32_t)pc * PB;
                cp_async_small_zfill<PB>(sb + off, kbase + pr * p.nbk1 + PB * pc, in ? (uint32_t)PB : 0u);
                cp_async_small_zfill<PB>(sb + RCH + off, vbase + pr * p.nbv1 + PB * pc, in ? (uint32_t)PB : 0u);
            }
        } else {
#pragma unroll
            for (int i = 0; i < CH * PPR / 32; ++i) {
                const int q = 32 * i + lane, row = q / PPR, pc = q % PPR;
                const bool in = pos0 + row < p.n_kv;
                const int64_t pr = in ? pos0 + row : 0;
                const uint32_t off = (uint32_t)row * ROWB + (uint32_t)((pc ^ (row & 7)) << 4);
                cp_async16_zfill(sb + off, kbase + pr * p.nbk1 + 16 * pc, in ? 16u : 0u);
                cp_async16_zfill(sb + RCH + off, vbase + pr * p.nbv1 + 16 * pc, in ? 16u : 0u);
            }
        }
        // The mask values of the chunk travel with it: read with ordinary loads at the point of use they were 34 % of the kernel's
        // stall samples (one L2 round trip per chunk in a warp that has nothing else to run; profiles/r02_fattn_mma_32k_ncu.md).
        if (p.mask) {
#pragma unroll
            for (int i = 0; i < 2; ++i) {
                const int q = 32 * i + lane, g = q >> 2, pc = q & 3;
                const int gc = min(g0 + g, ngq - 1), pos = pos0 + 8 * pc;
                const int valid = max(0, min(16, 2 * (p.n_kv - pos)));
                const char * src = valid > 0 ? p.mask + (int64_t)(gc / p.rk2) * p.nbm1 + 2 * (int64_t)pos : p.mask;
                cp_async16_zfill(sb + 2 * RCH + g * FMMA_MROW + 16 * pc, src, (uint32_t)valid);
            }
        }
    };
#pragma unroll
    for (int c = 0; c < NS - 1; ++c) {
        if (c < nch) issue(c);
        asm volatile("cp.async.commit_group;" ::: "memory");
    }
    __syncthreads();                                       // Q is in shared memory
    uint32_t qa[KS][4];
#pragma unroll
    for (int kk = 0; kk < KS; ++kk) {
        const int mi = lane >> 3, row = ((mi & 1) << 3) + (lane & 7), pc = 2 * kk + (mi >> 1);
        ldsm_x4(qs + row * ROWB + ((pc ^ (row & 7)) << 4), qa[kk]);
    }
    // per-row constants of this thread's two rows
    float slope[2];
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        const int gc = min(g0 + r + 8 * h, ngq - 1);
        const unsigned hd = (unsigned)(ikv * p.rk2 + gc % p.rk2);
        slope[h] = p.max_bias > 0.0f ? (hd < p.n_head_log2 ? powf(p.m0, (float)(hd + 1)) : powf(p.m1, (float)(2 * (hd - p.n_head_log2) + 1))) : 1.0f;
    }
    const bool two = g0 + 8 < ngq;                         // rows 8..15 hold queries (CTA-uniform)

    float M[2] = {-INFINITY, -INFINITY}, L[2] = {0.0f, 0.0f};
    float o[NB][4];
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) { o[nb][0] = 0.0f; o[nb][1] = 0.0f; o[nb][2] = 0.0f; o[nb][3] = 0.0f; }

    for (int c = 0; c < nch; ++c) {
        if (c + NS - 1 < nch) issue(c + NS - 1);
        asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group %0;" :: "n"(NS - 1) : "memory");
        __syncwarp();
        const uint32_t sraw = wb + (uint32_t)(c % NS) * STB, smk = sraw + 2 * RCH;
        const uint32_t sk = Q8 ? ft : sraw, sv = sk + CHB;
        const int pos0 = CH * (t0 + warp + NW * c);
        if constexpr (Q8) {                                 // raw blocks -> this warp's f16 tiles (same swizzle as the f16 stages)
            // Lane = one 32-value block of one row per pass.  A block starts 4-byte aligned (even blocks) or 2 bytes past (odd): nine
            // aligned words cover {d, q[32]} either way, one byte permute per word puts the codes in place (word-wise instead of
            // 2-byte loads: 89 instead of 120 instructions per block).
            constexpr int NBLK = D / 32;
            const __half2 c1152 = __half2(__ushort_as_half((unsigned short)0x6480), __ushort_as_half((unsigned short)0x6480));
#pragma unroll 2
            for (int i = 0; i < CH * NBLK / 32; ++i) {
                const int u = 32 * i + lane, row = u / NBLK, b = u % NBLK;
#pragma unroll
                for (int kv = 0; kv < 2; ++kv) {
                    const uint32_t blk = sraw + kv * RCH + (uint32_t)row * RAWB + 34 * b;
                    const bool odd = (blk & 2) != 0;
                    const uint32_t base = blk & ~3u, sel = odd ? 0x7654u : 0x5432u;
                    uint32_t w[9];
#pragma unroll
                    for (int k = 0; k < 9; ++k) w[k] = lds32(base + 4 * k);
                    const uint32_t db = odd ? (w[0] >> 16) : (w[0] & 0xffffu);
                    const __half dh = __ushort_as_half((unsigned short)db);
                    const __half2 d2 = __half2(dh, dh);
                    uint32_t o[16];
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const uint32_t x = __byte_perm(w[k], w[k + 1], sel) ^ 0x80808080u;      // four codes, biased to 0 .. 255
                        const uint32_t lo = __byte_perm(x, 0x64646464u, 0x4140), hi = __byte_perm(x, 0x64646464u, 0x4342);
                        const __half2 a0 = __hmul2(__hsub2(*reinterpret_cast<const __half2 *>(&lo), c1152), d2);
                        const __half2 a1 = __hmul2(__hsub2(*reinterpret_cast<const __half2 *>(&hi), c1152), d2);
                        o[2 * k] = *reinterpret_cast<const uint32_t *>(&a0); o[2 * k + 1] = *reinterpret_cast<const uint32_t *>(&a1);
                    }
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const int pc = 4 * b + j;
                        sts128_u4(sk + kv * CHB + (uint32_t)row * ROWB + (uint32_t)((pc ^ (row & 7)) << 4), make_uint4(o[4 * j], o[4 * j + 1], o[4 * j + 2], o[4 * j + 3]));
                    }
                }
            }
            __syncwarp();
        }
        float s[4][4];
#pragma unroll
        for (int j = 0; j < 4; ++j) { s[j][0] = 0.0f; s[j][1] = 0.0f; s[j][2] = 0.0f; s[j][3] = 0.0f; }
#pragma unroll
        for (int kk = 0; kk < KS; kk += 2) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                uint32_t kb[4];
                const int row = 8 * j + (lane & 7), pc = 2 * kk + (lane >> 3);
                ldsm_x4(sk + row * ROWB + ((pc ^ (row & 7)) << 4), kb);
                mma_f16(s[j], qa[kk], kb[0], kb[1]);
                mma_f16(s[j], qa[kk + 1], kb[2], kb[3]);
            }
        }
        // online softmax on the fragments
        float corr[2];
        uint32_t pa[2][4];
        bool moved = false;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            float x[8];
            float mx = -INFINITY;
            float2 mk[4];                                  // raw mask values of the row's 8 positions (the ALiBi slope is applied at use)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                mk[j] = make_float2(0.0f, 0.0f);
                if (p.mask) {
                    const uint32_t w = lds32(smk + (r + 8 * h) * FMMA_MROW + 16 * j + 4 * t);
                    mk[j] = __half22float2(*reinterpret_cast<const __half2 *>(&w));
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                    const int ic = pos0 + 8 * j + 2 * t + e;
                    float v = s[j][2 * h + e] * p.scale;
                    if (p.softcap != 0.0f) v = p.softcap * tanhf(v);
                    v += slope[h] * (e ? mk[j].y : mk[j].x);
                    if (ic >= p.n_kv || !(h == 0 || two)) v = -INFINITY;
                    x[2 * j + e] = v;
                    mx = fmaxf(mx, v);
                }
            }
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
            const float Mn = fmaxf(M[h], mx);
            corr[h] = Mn == -INFINITY ? 1.0f : __expf(M[h] - Mn);      // M = -inf: exp(-inf) = 0 scales the (zero) state
            float ps = 0.0f;
#pragma unroll
            for (int i = 0; i < 8; ++i) { x[i] = Mn == -INFINITY ? 0.0f : __expf(x[i] - Mn); ps += x[i]; }
            L[h] = L[h] * corr[h] + ps;
            M[h] = Mn;
            moved = moved || corr[h] != 1.0f;
            // P as the A operand of the second product: k-step kk2 takes n-blocks 2 kk2 (k 0..7) and 2 kk2 + 1 (k 8..15)
            pa[0][h] = pack_h2(x[0], x[1]); pa[0][2 + h] = pack_h2(x[2], x[3]);
            pa[1][h] = pack_h2(x[4], x[5]); pa[1][2 + h] = pack_h2(x[6], x[7]);
        }
        if (__any_sync(0xffffffffu, moved)) {
#pragma unroll
            for (int nb = 0; nb < NB; ++nb) { o[nb][0] *= corr[0]; o[nb][1] *= corr[0]; o[nb][2] *= corr[1]; o[nb][3] *= corr[1]; }
        }
#pragma unroll
        for (int kk2 = 0; kk2 < 2; ++kk2) {
#pragma unroll
            for (int nb = 0; nb < NB; nb += 2) {
                uint32_t vb[4];
                const int mi = lane >> 3, row = 16 * kk2 + ((mi & 1) << 3) + (lane & 7), pc = nb + (mi >> 1);
                ldsm_x4_t(sv + row * ROWB + ((pc ^ (row & 7)) << 4), vb);
                mma_f16(o[nb], pa[kk2], vb[0], vb[1]);
                mma_f16(o[nb + 1], pa[kk2], vb[2], vb[3]);
            }
        }
        __syncwarp();                                      // the stage is free for the copy issued in the next iteration
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");
#pragma unroll
    for (int h = 0; h < 2; ++h) {
        L[h] += __shfl_xor_sync(0xffffffffu, L[h], 1);
        L[h] += __shfl_xor_sync(0xffffffffu, L[h], 2);
    }
    __syncthreads();                                       // the merge buffers take the place of the stages (and of Q)
    float (*wacc)[16][D] = reinterpret_cast<float (*)[16][D]>(fm_smem);
    if (t == 0) { wm[warp][r] = M[0]; wl[warp][r] = L[0]; wm[warp][r + 8] = M[1]; wl[warp][r + 8] = L[1]; }
#pragma unroll
    for (int nb = 0; nb < NB; ++nb) {
        *reinterpret_cast<float2 *>(&wacc[warp][r][8 * nb + 2 * t]) = make_float2(o[nb][0], o[nb][1]);
        *reinterpret_cast<float2 *>(&wacc[warp][r + 8][8 * nb + 2 * t]) = make_float2(o[nb][2], o[nb][3]);
    }
    __syncthreads();
    for (int i = tid; i < 16 * D; i += 32 * NW) {
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

}  // namespace b200
