// fattn_tc.cu -- GGML_OP_FLASH_ATTN_EXT for prefill batches on the sm_100a tensor cores (f16 KV cache).
//
// Replaces the reference's tile / mma flash-attention kernels (ggml-cuda/fattn.cu:271-346 dispatch,
// fattn-mma-f16.cuh, fattn-tile-f16.cu) for batches of more than 8 query rows; semantics follow
// ggml_compute_forward_flash_attn_ext_f16 (ggml-cpu/ops.cpp:7000-7250): Q is rounded to f16 as the CPU path's
// q_to_vec_dot does, scores are f32, softmax is online, P and V are f16 with f32 accumulation.
//
// One CTA = 128 query rows of one head.  Per 128-position KV tile:
//   S = Q K^T     tcgen05.mma, Q and K (TMA, 128-byte swizzle) from shared memory, S (f32) into TMEM, double buffered
//   softmax       4 warps, thread = query row = TMEM lane: tcgen05.ld the row, scale / softcap / mask, online max and
//                 sum, P rounded to f16 and written back to TMEM with tcgen05.st; O (TMEM) is rescaled when the
//                 running max moved
//   O += P V      tcgen05.mma with A = P from TMEM and B = the V tile as it lies in the cache (rows = positions:
//                 an MN-major shared-memory operand)
// KV tiles whose mask block is entirely -inf (the causal upper triangle, other sequences) are found by a scan of
// the mask before the loop and never loaded.
#include "tc.cuh"
#include "../../include/b200_ops.h"

#include <cmath>

namespace b200 {

struct fat_params {
    const char * q; const char * mask; char * dst;
    int64_t nbq1, nbq2, nbq3, nbm1, nbd1;
    int N, n_head, n_kv, rk2, rk3, mask_vec;
    float scale, softcap, max_bias, m0, m1;
    unsigned n_head_log2;
    const uint8_t * flags;       // [q tiles][kv tiles] from mask_classify_kernel: 0 nothing to attend, 1 mixed, 2 mask is all zero
    uint32_t * xb;               // optional: the result also as bf16 rows [token][n_head * D] (the activations of the output projection's MMQ)
};

// ---- one pass over the mask for all heads: which 128 x 128 tiles are empty, which need no mask at all ----
// (the attention kernel used to scan its 128 mask rows per CTA, once per head, with one thread per row: at 32k context
// that scan and the per-element mask loads were most of the kernel's time)
__global__ void __launch_bounds__(256) mask_classify_kernel(const char * __restrict__ mask, int64_t nbm1, int N, int n_kv, int nt, int vec,
                                                            uint8_t * __restrict__ flags) {
    pdl_trigger(); pdl_wait();
    const int j = blockIdx.x, qt = blockIdx.y;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int lim = min(128, n_kv - 128 * j);
    int any = 0, nz = 0;
    if (vec && lim == 128) {
        // all 16 rows of the warp in flight at once (a row-by-row loop paid one memory round trip per row: 16 in series)
        uint2 u[16];
#pragma unroll
        for (int it = 0; it < 16; ++it) {
            const int qi = 128 * qt + warp + 8 * it;
            if (qi < N) u[it] = __ldg(reinterpret_cast<const uint2 *>(reinterpret_cast<const __half *>(mask + (int64_t)qi * nbm1) + 128 * j) + lane);
        }
#pragma unroll
        for (int it = 0; it < 16; ++it) {
            const int qi = 128 * qt + warp + 8 * it;
            if (qi < N) {
                any |= (u[it].x != 0xFC00FC00u) | (u[it].y != 0xFC00FC00u);
                nz |= ((u[it].x | u[it].y) & 0x7FFF7FFFu) != 0u;
            }
        }
    } else {
        for (int r = warp; r < 128; r += 8) {
            const int qi = 128 * qt + r;
            if (qi >= N) break;
            const __half * row = reinterpret_cast<const __half *>(mask + (int64_t)qi * nbm1) + 128 * j;
            for (int i = lane; i < lim; i += 32) {
                const unsigned short h = __half_as_ushort(row[i]);
                any |= h != 0xFC00u; nz |= (h & 0x7FFFu) != 0;
            }
        }
    }
    any = __syncthreads_or(any);
    nz = __syncthreads_or(nz);
    if (threadIdx.x == 0) flags[(int64_t)qt * nt + j] = !any ? 0 : ((nz || lim < 128) ? 1 : 2);
}


constexpr int FAT_THREADS = 192;
constexpr int FAT_MAX_TILES = 512;

__device__ __forceinline__ float ex2(float x) { float y; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x)); return y; }
__device__ __forceinline__ uint32_t pack_f16(float lo, float hi) {
    uint32_t d; asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo)); return d;
}

template <int D>
__global__ void __launch_bounds__(FAT_THREADS, 1)
fattn_tc_kernel(const __grid_constant__ CUtensorMap kmap, const __grid_constant__ CUtensorMap vmap, const fat_params p) {
    constexpr int NA = D / 64;                   // 64-element (128-byte) swizzle atoms per row of Q / K / V
    constexpr int ATOM = 128 * 128;              // 128 rows x 128 bytes
    constexpr int TILE = NA * ATOM;
    constexpr float LOG2E = 1.4426950408889634f;
    extern __shared__ uint8_t smem_dyn[];
    const uint32_t base = (smem_u32(smem_dyn) + 1023u) & ~1023u;
    const uint32_t sQ = base, sK = sQ + TILE, sV = sK + 2 * TILE, sBar = sV + 2 * TILE;
    const uint32_t k_full = sBar, k_empty = sBar + 16, v_full = sBar + 32, v_empty = sBar + 48, s_full = sBar + 64, s_empty = sBar + 80,
                   p_full = sBar + 96, o_done = sBar + 104, tmem_slot = sBar + 112;
    uint8_t * gen = smem_dyn + (base - smem_u32(smem_dyn));
    int * act = reinterpret_cast<int *>(gen + 5 * TILE + 128);                 // [FAT_MAX_TILES] any unmasked entry in KV tile j
    unsigned short * tiles = reinterpret_cast<unsigned short *>(act + FAT_MAX_TILES);   // compacted list of active tiles
    int * nact_p = reinterpret_cast<int *>(tiles + FAT_MAX_TILES);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int q0 = blockIdx.x * 128, iq2 = blockIdx.y % p.n_head, iq3 = blockIdx.y / p.n_head;
    const int nt = (p.n_kv + 127) / 128;

    if (threadIdx.x == 0) {
        for (int i = 0; i < 2; ++i) {
            mbar_init(k_full + 8 * i, 1); mbar_init(k_empty + 8 * i, 1); mbar_init(v_full + 8 * i, 1); mbar_init(v_empty + 8 * i, 1);
            mbar_init(s_full + 8 * i, 1); mbar_init(s_empty + 8 * i, 4);
        }
        mbar_init(p_full, 4); mbar_init(o_done, 1);
        mbar_fence_init();
        tma_prefetch_desc(&kmap); tma_prefetch_desc(&vmap);
    }
    if (warp == 1) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
    for (int j = threadIdx.x; j < nt; j += FAT_THREADS) act[j] = p.mask ? 0 : (128 * (j + 1) <= p.n_kv ? 2 : 1);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = lds32(tmem_slot);
    pdl_trigger();
    pdl_wait();                                  // Q, the KV cache rows and the mask come from preceding kernels

    const int row = (warp & 3) * 32 + lane;      // softmax warps 2..5 own TMEM lane quarters 2,3,0,1
    const int qi = q0 + row;
    const bool valid = warp >= 2 && qi < p.N;
    const __half * mrow = p.mask && valid ? reinterpret_cast<const __half *>(p.mask + (int64_t)qi * p.nbm1) : nullptr;

    if (warp >= 2) {
        // ---- Q: f32 -> f16 into the swizzled K-major A tile (rows past N are zero) ----
        const float * qrow = reinterpret_cast<const float *>(p.q + (int64_t)qi * p.nbq1 + (int64_t)iq2 * p.nbq2 + (int64_t)iq3 * p.nbq3);
#pragma unroll
        for (int u = 0; u < D / 8; ++u) {
            uint32_t w[4] = {0, 0, 0, 0};
            if (valid) {
                const float4 a = *reinterpret_cast<const float4 *>(qrow + 8 * u), b = *reinterpret_cast<const float4 *>(qrow + 8 * u + 4);
                w[0] = pack_f16(a.x, a.y); w[1] = pack_f16(a.z, a.w); w[2] = pack_f16(b.x, b.y); w[3] = pack_f16(b.z, b.w);
            }
            sts128(sQ + (u >> 3) * ATOM + row * 128 + (((u & 7) ^ (row & 7)) << 4), w[0], w[1], w[2], w[3]);
        }
        fence_proxy_async();
        ktrace_event_at(2, 64, gridDim.x - 1);
        // ---- which KV tiles have anything to attend to? ----
        if (p.mask && p.flags) {
            for (int j = threadIdx.x - 64; j < nt; j += FAT_THREADS - 64) act[j] = p.flags[(int64_t)blockIdx.x * nt + j];
        } else if (p.mask) {
            for (int j = 0; j < nt; ++j) {
                bool any = false;
                if (mrow) {
                    const int lim = min(128, p.n_kv - 128 * j);
                    if (p.mask_vec && lim == 128) {
                        const uint4 * mp = reinterpret_cast<const uint4 *>(mrow + 128 * j);
#pragma unroll 4
                        for (int g = 0; g < 16; ++g) {
                            const uint4 v = __ldg(mp + g);
                            any |= (v.x != 0xFC00FC00u) | (v.y != 0xFC00FC00u) | (v.z != 0xFC00FC00u) | (v.w != 0xFC00FC00u);
                        }
                    } else {
                        for (int i = 0; i < lim; ++i) any |= __half_as_ushort(mrow[128 * j + i]) != 0xFC00u;
                    }
                }
                if (__any_sync(0xffffffffu, any) && lane == 0) atomicOr(&act[j], 1);
            }
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int n = 0;
        for (int j = 0; j < nt; ++j) if (act[j]) tiles[n++] = (unsigned short)(j | (act[j] == 2 ? 0x8000 : 0));      // bit 15: no mask needed
        *nact_p = n;
    }
    __syncthreads();
    const int nact = *nact_p;
    ktrace_event_at(3, 64, gridDim.x - 1);

    if (warp == 0) {
        if (lane == 0) {                         // TMA producer: K and V tiles of the active KV tiles, two stages each
            const int ik2 = iq2 / p.rk2, ik3 = iq3 / p.rk3;
            for (int jj = 0; jj < nact; ++jj) {
                const int j = tiles[jj] & 0x7FFF, st = jj & 1;
                const uint32_t ph = ((jj >> 1) & 1) ^ 1;
                mbar_wait_b(k_empty + 8 * st, ph);
                mbar_expect_tx(k_full + 8 * st, TILE);
#pragma unroll
                for (int a = 0; a < NA; ++a) tma_load_4d(sK + st * TILE + a * ATOM, &kmap, 64 * a, 128 * j, ik2, ik3, k_full + 8 * st);
                mbar_wait_b(v_empty + 8 * st, ph);
                mbar_expect_tx(v_full + 8 * st, TILE);
#pragma unroll
                for (int a = 0; a < NA; ++a) tma_load_4d(sV + st * TILE + a * ATOM, &vmap, 64 * a, 128 * j, ik2, ik3, v_full + 8 * st);
            }
        }
    } else if (warp == 1) {
        if (nact > 0 && elect_one()) {           // tensor-core issue
            const uint32_t idesc_s = umma_idesc(0, 128, 0), idesc_o = umma_idesc(0, D, 1);
            auto issue_s = [&](int jj) {
                const int st = jj & 1;
                mbar_wait_b(k_full + 8 * st, (jj >> 1) & 1);
                mbar_wait_b(s_empty + 8 * st, ((jj >> 1) & 1) ^ 1);       // softmax has read the previous S in this buffer
                tc_fence_after();
#pragma unroll
                for (int kk = 0; kk < D / 16; ++kk) {
                    const uint32_t off = (kk >> 2) * ATOM + (kk & 3) * 32;
                    tc_mma_f16(tmem_base + 128 * st, umma_desc_sw128(sQ + off), umma_desc_sw128(sK + st * TILE + off), idesc_s, kk != 0);
                }
                tc_commit(s_full + 8 * st);
                tc_commit(k_empty + 8 * st);
            };
            issue_s(0);
            for (int jj = 0; jj < nact; ++jj) {
                if (jj + 1 < nact) issue_s(jj + 1);
                const int st = jj & 1;
                mbar_wait_b(p_full, jj & 1);
                mbar_wait_b(v_full + 8 * st, (jj >> 1) & 1);
                tc_fence_after();
#pragma unroll
                for (int kk = 0; kk < 8; ++kk)   // 16 KV positions per instruction: 8 TMEM columns of P, 16 rows (2048 B) of V
                    tc_mma_f16_ts(tmem_base + 256, tmem_base + 384 + 8 * kk, umma_desc_sw128_mn(sV + st * TILE + kk * 2048, ATOM), idesc_o,
                                  (uint32_t)((jj | kk) != 0));
                tc_commit(o_done);
                tc_commit(v_empty + 8 * st);
            }
        }
    } else {
        // ================= softmax warps =================
        const uint32_t lane_base = tmem_base + ((uint32_t)((warp & 3) * 32) << 16);
        const unsigned h = iq2;
        const float slope = p.max_bias > 0.0f ? (h < p.n_head_log2 ? powf(p.m0, (float)(h + 1)) : powf(p.m1, (float)(2 * (h - p.n_head_log2) + 1))) : 1.0f;
        const float sc2 = p.softcap != 0.0f ? p.scale : p.scale * LOG2E, ms2 = slope * LOG2E;
        float m = -INFINITY, l = 0.0f;           // running max (log2 domain) and sum
        for (int jj = 0; jj < nact; ++jj) {
            const int j = tiles[jj] & 0x7FFF, st = jj & 1;
            const bool nomask = (tiles[jj] & 0x8000) != 0;
            // a tile that needs its mask: the 16 loads of this thread's mask row go out before the wait for S (they do not depend on it);
            // issued one per use inside the loop below they were 16 memory round trips in series, 5 us of a 17 us CTA
            const int lim = p.n_kv - 128 * j;
            const bool mvec = !nomask && mrow && p.mask_vec && lim >= 128;
            uint4 mreg[16];
            if (mvec) {
#pragma unroll
                for (int g = 0; g < 16; ++g) mreg[g] = __ldg(reinterpret_cast<const uint4 *>(mrow + 128 * j) + g);
            }
            mbar_wait_b(s_full + 8 * st, (jj >> 1) & 1);
            tc_fence_after();
            ktrace_event_at(4, 64, gridDim.x - 1);
            float s[128];
#pragma unroll
            for (int c = 0; c < 4; ++c) {
                uint32_t v[32];
                tmem_ld32(lane_base + 128 * st + 32 * c, v);
                tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) s[32 * c + i] = __uint_as_float(v[i]);
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(s_empty + 8 * st);
            ktrace_event_at(5, 64, gridDim.x - 1);
            // scale, softcap, mask (all in the log2 domain), positions past n_kv and rows past N are -inf
            float rmax = -INFINITY;
            if (nomask) {                        // a whole tile, every row attends to every position: scale (and softcap) only
#pragma unroll
                for (int i = 0; i < 128; ++i) {
                    float x = s[i] * sc2;
                    if (p.softcap != 0.0f) x = p.softcap * LOG2E * tanhf(x);
                    x = valid ? x : -INFINITY;
                    s[i] = x;
                    rmax = fmaxf(rmax, x);
                }
            } else
#pragma unroll
            for (int g = 0; g < 16; ++g) {
                float mv[8];
                if (mvec) {
                    const uint4 u = mreg[g];
                    const float2 a = __half22float2(*reinterpret_cast<const __half2 *>(&u.x)), b = __half22float2(*reinterpret_cast<const __half2 *>(&u.y));
                    const float2 c = __half22float2(*reinterpret_cast<const __half2 *>(&u.z)), d = __half22float2(*reinterpret_cast<const __half2 *>(&u.w));
                    mv[0] = a.x; mv[1] = a.y; mv[2] = b.x; mv[3] = b.y; mv[4] = c.x; mv[5] = c.y; mv[6] = d.x; mv[7] = d.y;
                } else {
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const int i = 8 * g + e;
                        mv[e] = i < lim ? (mrow ? __half2float(mrow[128 * j + i]) : 0.0f) : -INFINITY;
                    }
                }
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    float x = s[8 * g + e] * sc2;
                    if (p.softcap != 0.0f) x = p.softcap * LOG2E * tanhf(x);
                    x = valid ? fmaf(mv[e], ms2, x) : -INFINITY;
                    s[8 * g + e] = x;
                    rmax = fmaxf(rmax, x);
                }
            }
            const float mn = fmaxf(m, rmax);
            const float mref = mn == -INFINITY ? 0.0f : mn;       // fully masked so far: every p is 0
            const float alpha = ex2(m - mref);                     // m = -inf -> 0
            float sum = 0.0f;
            uint32_t pk[64];
#pragma unroll
            for (int i = 0; i < 64; ++i) {
                const float a = ex2(s[2 * i] - mref), b = ex2(s[2 * i + 1] - mref);
                pk[i] = pack_f16(a, b);
                // the sum uses the rounded values so that P / l is normalised over what the MMA really adds up
                const float2 r = __half22float2(*reinterpret_cast<const __half2 *>(&pk[i]));
                sum += r.x + r.y;
            }
            l = l * alpha + sum;
            m = mn;
            ktrace_event_at(6, 64, gridDim.x - 1);
            if (jj > 0) { mbar_wait_b(o_done, (jj - 1) & 1); tc_fence_after(); }       // previous P V finished: P is free, O is complete
            ktrace_event_at(7, 64, gridDim.x - 1);
            {
                uint32_t (&lo)[32] = *reinterpret_cast<uint32_t (*)[32]>(&pk[0]);
                uint32_t (&hi)[32] = *reinterpret_cast<uint32_t (*)[32]>(&pk[32]);
                tmem_st32(lane_base + 384, lo);
                tmem_st32(lane_base + 416, hi);
            }
            if (jj > 0 && !__all_sync(0xffffffffu, alpha == 1.0f)) {
#pragma unroll
                for (int c = 0; c < D / 32; ++c) {
                    uint32_t v[32];
                    tmem_ld32(lane_base + 256 + 32 * c, v);
                    tmem_ld_wait();
#pragma unroll
                    for (int i = 0; i < 32; ++i) v[i] = __float_as_uint(__uint_as_float(v[i]) * alpha);
                    tmem_st32(lane_base + 256 + 32 * c, v);
                }
            }
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_full);
            ktrace_event_at(8, 64, gridDim.x - 1);
        }
        // ---- epilogue: O / l -> dst[(iq3*N + q)*n_head + head][D] ----
        float * out = reinterpret_cast<float *>(p.dst + (((int64_t)iq3 * p.N + qi) * p.n_head + iq2) * p.nbd1);
        if (nact > 0) { mbar_wait_b(o_done, (nact - 1) & 1); tc_fence_after(); }
        ktrace_event_at(9, 64, gridDim.x - 1);
        const float inv = l > 0.0f ? 1.0f / l : 0.0f;
#pragma unroll
        for (int c = 0; c < D / 32; ++c) {
            uint32_t v[32];
            if (nact > 0) { tmem_ld32(lane_base + 256 + 32 * c, v); tmem_ld_wait(); }
            if (valid) {
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    float4 o;
                    o.x = nact > 0 ? __uint_as_float(v[i]) * inv : 0.0f; o.y = nact > 0 ? __uint_as_float(v[i + 1]) * inv : 0.0f;
                    o.z = nact > 0 ? __uint_as_float(v[i + 2]) * inv : 0.0f; o.w = nact > 0 ? __uint_as_float(v[i + 3]) * inv : 0.0f;
                    *reinterpret_cast<float4 *>(out + 32 * c + i) = o;
                    if (p.xb) {                  // same values, rounded as x_to_bf16_kernel rounds them
                        uint2 w;                 // cvt.rn.bf16x2.f32 {hi, lo}, as pack_bf16 of mmq.cu
                        asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w.x) : "f"(o.y), "f"(o.x));
                        asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w.y) : "f"(o.w), "f"(o.z));
                        *reinterpret_cast<uint2 *>(p.xb + ((((int64_t)iq3 * p.N + qi) * p.n_head + iq2) * D + 32 * c + i) / 2) = w;
                    }
                }
            }
        }
    }
    ktrace_event_at(10, 64, gridDim.x - 1);
    tc_fence_before();
    __syncthreads();
    if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, 512); }
}

template <int D> static size_t fat_smem() { return 1024 + 5 * (size_t)(D / 64) * 128 * 128 + 128 + FAT_MAX_TILES * 4 + FAT_MAX_TILES * 2 + 64; }

template <int D>
static int fat_launch(const CUtensorMap & kmap, const CUtensorMap & vmap, const fat_params & p, dim3 grid, cudaStream_t st) {
    static bool configured[16] = {false};   // per device
    const size_t smem = fat_smem<D>();
    int dev = 0;
    B200_CUDA_CHECK(cudaGetDevice(&dev)); dev &= 15;
    if (!configured[dev]) {
        B200_CUDA_CHECK(cudaFuncSetAttribute(fattn_tc_kernel<D>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        configured[dev] = true;
    }
    cudaLaunchConfig_t lc = {};
    lc.gridDim = grid; lc.blockDim = dim3(FAT_THREADS); lc.dynamicSmemBytes = smem; lc.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = pdl_default() ? 1 : 0;
    B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, fattn_tc_kernel<D>, kmap, vmap, p));
    ++g_launches;
    return 0;
}

static int fat_min_rows() { static const int v = [] { const char * e = getenv("B200_FATTN_TC_MIN_ROWS"); return e ? atoi(e) : 9; }(); return v; }

// 1 = launched, 0 = not applicable (the caller uses the vector kernel), < 0 / > 0 = error
int64_t fattn_tc_flag_bytes(const b200_tensor * q, const b200_tensor * k) {
    if (q->ne[1] < fat_min_rows() || k->ne[1] > 128 * FAT_MAX_TILES || (k->type != T_F16 && k->type != T_Q8_0)) return 0;
    return ((((q->ne[1] + 127) / 128) * ((k->ne[1] + 127) / 128) + 255) & ~(int64_t)255) + 256;
}
int fattn_tc_try(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask, const b200_tensor * dst,
                 float scale, float max_bias, float logit_softcap, void * scratch, int64_t scratch_bytes, cudaStream_t st, int * rc, void * xb) {
    *rc = 0;
    const int D = (int)q->ne[0];
    if (k->type != T_F16 || v->type != T_F16 || (D != 64 && D != 128) || v->ne[0] != D) return 0;
    if (q->ne[1] < fat_min_rows() || k->ne[1] > 128 * FAT_MAX_TILES || !encode_tiled()) return 0;
    if (((uintptr_t)k->data | (uintptr_t)v->data | k->nb[1] | k->nb[2] | k->nb[3] | v->nb[1] | v->nb[2] | v->nb[3]) & 15) return 0;
    if (((uintptr_t)q->data | q->nb[1] | q->nb[2] | q->nb[3] | (uintptr_t)dst->data | dst->nb[1]) & 15) return 0;
    if (k->nb[0] != 2 || v->nb[0] != 2) return 0;
    CUtensorMap kmap, vmap;
    for (int which = 0; which < 2; ++which) {
        const b200_tensor * t = which ? v : k;
        const cuuint64_t dims[4] = {(cuuint64_t)D, (cuuint64_t)t->ne[1], (cuuint64_t)t->ne[2], (cuuint64_t)t->ne[3]};
        const cuuint64_t strides[3] = {(cuuint64_t)t->nb[1], (cuuint64_t)t->nb[2], (cuuint64_t)t->nb[3]};
        const cuuint32_t box[4] = {64, 128, 1, 1};
        const cuuint32_t es[4] = {1, 1, 1, 1};
        if (encode_tiled()(which ? &vmap : &kmap, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, t->data, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) return 0;
    }
    fat_params p = {};
    p.q = (const char *)q->data; p.mask = mask ? (const char *)mask->data : nullptr; p.dst = (char *)dst->data;
    p.xb = (uint32_t *)xb;
    p.nbq1 = q->nb[1]; p.nbq2 = q->nb[2]; p.nbq3 = q->nb[3]; p.nbm1 = mask ? mask->nb[1] : 0; p.nbd1 = dst->nb[1];
    p.N = (int)q->ne[1]; p.n_head = (int)q->ne[2]; p.n_kv = (int)k->ne[1];
    p.rk2 = (int)(q->ne[2] / k->ne[2]); p.rk3 = (int)(q->ne[3] / k->ne[3]);
    p.mask_vec = mask && (((uintptr_t)mask->data | mask->nb[1]) & 15) == 0;
    if (logit_softcap != 0.0f) scale /= logit_softcap;
    p.scale = scale; p.softcap = logit_softcap; p.max_bias = max_bias;
    p.n_head_log2 = 1u << (unsigned)floorf(log2f((float)p.n_head));
    p.m0 = powf(2.0f, -max_bias / p.n_head_log2); p.m1 = powf(2.0f, -(max_bias / 2.0f) / p.n_head_log2);
    const dim3 grid((unsigned)((q->ne[1] + 127) / 128), (unsigned)(q->ne[2] * q->ne[3]));
    const int64_t fbytes = fattn_tc_flag_bytes(q, k);
    if (mask && scratch && fbytes > 0 && scratch_bytes >= fbytes) {
        uint8_t * flags = reinterpret_cast<uint8_t *>(((uintptr_t)scratch + 255) & ~(uintptr_t)255);
        const int nt = (int)((k->ne[1] + 127) / 128);
        cudaLaunchConfig_t lc = {};
        lc.gridDim = dim3((unsigned)nt, grid.x); lc.blockDim = dim3(256); lc.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        lc.attrs = at; lc.numAttrs = pdl_default() ? 1 : 0;
        if (cudaLaunchKernelEx(&lc, mask_classify_kernel, (const char *)mask->data, (int64_t)mask->nb[1], (int)q->ne[1], (int)k->ne[1], nt, (int)p.mask_vec, flags) != cudaSuccess) { *rc = -100; return 1; }
        ++g_launches;
        p.flags = flags;
    }
    *rc = D == 128 ? fat_launch<128>(kmap, vmap, p, grid, st) : fat_launch<64>(kmap, vmap, p, grid, st);
    return 1;
}

// ---- q8_0 KV cache -> f16 for the tensor-core kernel (prefill over a quantised cache) ----
// The reference dequantises K/V to f16 the same way before its tensor-core attention (ggml-cuda/fattn-common.cuh:700-760,
// launch_fattn with need_f16_K / need_f16_V).  One thread per 32-element block.
__global__ void __launch_bounds__(256) kv_q8_0_to_f16_kernel(const char * __restrict__ src, int64_t nb1, int64_t nb2, int64_t nb3,
                                                             __half * __restrict__ dst, int D, int n_kv, int n_head) {
    pdl_trigger(); pdl_wait();
    const int bpr = D >> 5;
    const int64_t idx = (int64_t)blockIdx.x * 256 + threadIdx.x;
    if (idx >= (int64_t)n_kv * bpr) return;
    const int pos = (int)(idx / bpr), b = (int)(idx % bpr);
    const int h = blockIdx.y % n_head, i3 = blockIdx.y / n_head;
    const unsigned short * blk = reinterpret_cast<const unsigned short *>(src + (int64_t)pos * nb1 + (int64_t)h * nb2 + (int64_t)i3 * nb3 + (int64_t)b * 34);
    const float d = __half2float(__ushort_as_half(blk[0]));
    uint32_t out[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const unsigned short w = blk[1 + i];
        const float lo = d * (float)(signed char)(w & 0xFF), hi = d * (float)(signed char)(w >> 8);
        out[i] = (uint32_t)__half_as_ushort(__float2half_rn(lo)) | ((uint32_t)__half_as_ushort(__float2half_rn(hi)) << 16);
    }
    uint4 * o = reinterpret_cast<uint4 *>(dst + ((((int64_t)i3 * n_head + h) * n_kv + pos) * D + b * 32));
#pragma unroll
    for (int i = 0; i < 4; ++i) o[i] = make_uint4(out[4 * i], out[4 * i + 1], out[4 * i + 2], out[4 * i + 3]);
}

int64_t fattn_tc_q8_scratch(const b200_tensor * q, const b200_tensor * k) {
    if (k->type != T_Q8_0 || q->ne[1] < fat_min_rows() || k->ne[1] > 128 * FAT_MAX_TILES || !encode_tiled()) return 0;
    return 2 * ((k->ne[0] * k->ne[1] * k->ne[2] * k->ne[3] * 2 + 255) & ~(int64_t)255) + 256 + fattn_tc_flag_bytes(q, k);
}

// q8_0 K and V: dequantise into scratch, then the f16 kernel.  Returns 1 when it ran (rc = its status).
int fattn_tc_try_q8(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask, const b200_tensor * dst,
                    float scale, float max_bias, float logit_softcap, void * scratch, int64_t scratch_bytes, cudaStream_t st, int * rc, void * xb) {
    *rc = 0;
    const int64_t need = fattn_tc_q8_scratch(q, k);
    if (need == 0 || v->type != T_Q8_0 || !scratch || scratch_bytes < need || v->ne[0] != k->ne[0] || v->ne[1] != k->ne[1] || v->ne[2] != k->ne[2] || v->ne[3] != k->ne[3]) return 0;
    if (((k->nb[1] | k->nb[2] | k->nb[3] | v->nb[1] | v->nb[2] | v->nb[3]) & 1) || ((uintptr_t)k->data & 1) || ((uintptr_t)v->data & 1)) return 0;
    const int D = (int)k->ne[0], n_kv = (int)k->ne[1], n_head = (int)k->ne[2];
    b200_tensor kv[2];
    for (int which = 0; which < 2; ++which) {
        const b200_tensor * t = which ? v : k;
        const int64_t fb = fattn_tc_flag_bytes(q, k);               // the mask flags sit in front
        __half * out = reinterpret_cast<__half *>((((uintptr_t)scratch + 255) & ~(uintptr_t)255) + (uintptr_t)fb + (uintptr_t)which * (uintptr_t)((need - fb - 256) / 2));
        cudaLaunchConfig_t lc = {};
        lc.gridDim = dim3((unsigned)(((int64_t)n_kv * (D / 32) + 255) / 256), (unsigned)(t->ne[2] * t->ne[3])); lc.blockDim = dim3(256); lc.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        lc.attrs = at; lc.numAttrs = pdl_default() ? 1 : 0;
        if (cudaLaunchKernelEx(&lc, kv_q8_0_to_f16_kernel, (const char *)t->data, t->nb[1], t->nb[2], t->nb[3], out, D, n_kv, n_head) != cudaSuccess) { *rc = -100; return 1; }
        ++g_launches;
        b200_tensor & f = kv[which];
        f.data = out; f.type = T_F16;
        f.ne[0] = D; f.ne[1] = n_kv; f.ne[2] = t->ne[2]; f.ne[3] = t->ne[3];
        f.nb[0] = 2; f.nb[1] = 2 * D; f.nb[2] = f.nb[1] * n_kv; f.nb[3] = f.nb[2] * t->ne[2];
    }
    return fattn_tc_try(q, &kv[0], &kv[1], mask, dst, scale, max_bias, logit_softcap, scratch, fattn_tc_flag_bytes(q, k), st, rc, xb);
}

}  // namespace b200
