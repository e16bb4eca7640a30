// common.cuh -- shared device helpers for the B200 (sm_100a) quantized forward-pass kernels.
// Block formats follow the GGUF spec the reference defines in ggml/src/ggml-common.h:167-326;
// written from the format description, not copied.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>

namespace b200 {

// ggml type ids (reference ggml/include/ggml.h:352-366)
enum : int { T_F32 = 0, T_F16 = 1, T_Q4_0 = 2, T_Q4_1 = 3, T_Q5_0 = 6, T_Q5_1 = 7, T_Q8_0 = 8, T_Q8_1 = 9, T_Q2_K = 10, T_Q3_K = 11, T_Q4_K = 12, T_Q5_K = 13, T_Q6_K = 14, T_IQ4_NL = 20, T_IQ4_XS = 23, T_I32 = 26, T_BF16 = 30 };

template <int TYPE> struct fmt;   // bytes per block, elements per block
template <> struct fmt<T_Q4_0> { static constexpr int BB = 18,  BE = 32;  };
template <> struct fmt<T_Q8_0> { static constexpr int BB = 34,  BE = 32;  };
template <> struct fmt<T_Q4_1> { static constexpr int BB = 20,  BE = 32;  };
template <> struct fmt<T_Q5_0> { static constexpr int BB = 22,  BE = 32;  };
template <> struct fmt<T_Q5_1> { static constexpr int BB = 24,  BE = 32;  };
template <> struct fmt<T_Q8_1> { static constexpr int BB = 36,  BE = 32;  };
template <> struct fmt<T_Q2_K> { static constexpr int BB = 84,  BE = 256; };
template <> struct fmt<T_Q3_K> { static constexpr int BB = 110, BE = 256; };
template <> struct fmt<T_Q4_K> { static constexpr int BB = 144, BE = 256; };
template <> struct fmt<T_IQ4_NL> { static constexpr int BB = 18,  BE = 32;  };
template <> struct fmt<T_IQ4_XS> { static constexpr int BB = 136, BE = 256; };
template <> struct fmt<T_Q5_K> { static constexpr int BB = 176, BE = 256; };
template <> struct fmt<T_Q6_K> { static constexpr int BB = 210, BE = 256; };

__host__ __device__ inline int block_bytes(int t) {
    switch (t) { case T_IQ4_NL: return 18; case T_IQ4_XS: return 136; case T_Q4_0: return 18; case T_Q4_1: return 20; case T_Q5_0: return 22; case T_Q5_1: return 24; case T_Q8_0: return 34; case T_Q8_1: return 36; case T_Q2_K: return 84; case T_Q3_K: return 110; case T_Q4_K: return 144;
                 case T_Q5_K: return 176; case T_Q6_K: return 210; case T_F32: return 4; case T_F16: return 2; case T_I32: return 4; case T_BF16: return 2; }
    return 0;
}
__host__ __device__ inline int block_elems(int t) {
    switch (t) { case T_IQ4_NL: case T_Q4_0: case T_Q4_1: case T_Q5_0: case T_Q5_1: case T_Q8_0: case T_Q8_1: return 32; case T_IQ4_XS: case T_Q2_K: case T_Q3_K: case T_Q4_K: case T_Q5_K: case T_Q6_K: return 256;
                 case T_F32: case T_F16: case T_I32: case T_BF16: return 1; }
    return 0;
}

// ---- shared-memory access at 2-byte alignment ------------------------------------------------
// GGUF blocks are only 2-byte aligned (18/34/210-byte strides).  A 32-bit value at a 2-byte
// aligned shared address = funnel shift of the two aligned words that cover it.
__device__ __forceinline__ uint32_t lds32(uint32_t saddr) {
    uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(saddr)); return v;
}
__device__ __forceinline__ uint32_t lds16(uint32_t saddr) {
    uint16_t v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(saddr)); return v;
}
__device__ __forceinline__ int4 lds128(uint32_t saddr) {
    int4 v; asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(saddr)); return v;
}
// N consecutive 32-bit words starting at a 2-byte aligned shared address
template <int N> __device__ __forceinline__ void lds_words_a2(uint32_t saddr, int (&out)[N]) {
    const uint32_t base = saddr & ~3u, sh = (saddr & 2u) * 8u;
    uint32_t w[N + 1];
#pragma unroll
    for (int i = 0; i <= N; ++i) w[i] = lds32(base + 4 * i);
#pragma unroll
    for (int i = 0; i < N; ++i) out[i] = (int)__funnelshift_r(w[i], w[i + 1], sh);
}

__device__ __forceinline__ float half_bits_to_float(uint32_t h16) {
    return __half2float(__ushort_as_half((unsigned short)h16));
}

// The 4-bit table formats (IQ4_NL, IQ4_XS; table: reference ggml-common.h:1077-1079): eight indices of a word -> the table's int8
// values of its four low nibbles (elements j .. j+3 of a block) and of its four high nibbles (elements j+16 ..).  PRMT picks any of
// eight bytes per selector nibble, so a 16-entry table is two picks (entries 0-7, entries 8-15, on index & 7) and a third that
// chooses between them on bit 3 of the index; two consecutive bytes of the word (four nibbles) are looked up at a time, and the
// results are de-interleaved into the low-nibble and the high-nibble word.
__device__ __forceinline__ void iq4_lookup8(uint32_t q, int & vlo, int & vhi) {
    constexpr uint32_t T0 = 0xbfad9881u, T1 = 0xf6eaddcfu, T2 = 0x26190d01u, T3 = 0x71594535u;      // -127 -104 -83 -65 | -49 -35 -22 -10 | 1 13 25 38 | 53 69 89 113
    uint32_t r[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const uint32_t idx = q >> (16 * i);                       // four indices: byte 2i low, high, byte 2i+1 low, high
        const uint32_t a = __byte_perm(T0, T1, idx & 0x7777u), b = __byte_perm(T2, T3, idx & 0x7777u);
        r[i] = __byte_perm(a, b, 0x3210u | ((idx & 0x8888u) >> 1));
    }
    vlo = (int)__byte_perm(r[0], r[1], 0x6420u);
    vhi = (int)__byte_perm(r[0], r[1], 0x7531u);
}

// ---- mbarrier + bulk async copy (TMA 1-D) ----------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t}" :: "r"(bar), "r"(parity) : "memory");
}
// global -> shared bulk copy, completion signalled on an mbarrier (SASS: UBLKCP).  src/dst 16-byte
// aligned, bytes a multiple of 16.
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void * src, uint32_t bytes, uint32_t bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}
// same, with an L2 evict-first policy: weights are streamed once per token
__device__ __forceinline__ void bulk_g2s_stream(uint32_t dst, const void * src, uint32_t bytes, uint32_t bar, uint64_t pol) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                 :: "r"(dst), "l"(src), "r"(bytes), "r"(bar), "l"(pol) : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
    uint64_t p; asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p)); return p;
}
// ---- kernel timeline trace (debug; b200_ktrace_enable / b200_ktrace_dump, ktrace.cu) ------------------------------
// Every kernel of the library enters through pdl_trigger() and passes pdl_wait() before it touches what the previous
// kernel wrote, so hooking these two gives a timeline of a whole decode step as it really runs (CUDA graph replay, PDL
// overlap) -- which ncu, serialising the launches, cannot show.  Thread 0 of block 0 appends
// (globaltimer, gridid, event, grid size, block size); off (null pointer in constant memory) it costs one predicate.
// Events 0 = entry, 1 = PDL wait passed; kernels may add their own phase marks (2..15).
// Compiled in only with -DB200_KTRACE (B200_KTRACE=1 python -m llamacog_b200.build): the pointer lives in a __constant__
// bank of its own, and the first touch of that bank by every kernel is a constant-cache miss -- measured at 0.4 us per
// launch (roofline 0.53 -> 0.49 on the 192 GEMVs of a token), too much for a hook that is normally off.
constexpr unsigned long long KTRACE_CAP = 1ull << 20;   // events
#ifndef B200_KTRACE
__device__ __forceinline__ void ktrace_event(int, int = 0) {}
__device__ __forceinline__ void ktrace_event_at(int, int, int) {}
#else
static __constant__ unsigned long long * c_ktrace;      // one copy per translation unit, all set by b200_ktrace_enable
__device__ __forceinline__ void ktrace_event(int ev, int thread = 0) {     // recorded by `thread` of block 0
    unsigned long long * t = c_ktrace;
    if (t != nullptr && ((threadIdx.x ^ thread) | blockIdx.x | blockIdx.y | blockIdx.z) == 0) {
        unsigned long long now, gid;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        asm volatile("mov.u64 %0, %%gridid;" : "=l"(gid));
        const unsigned long long slot = atomicAdd(t, 1ull);
        if (slot < KTRACE_CAP) {
            t[1 + 2 * slot] = now;
            t[2 + 2 * slot] = (gid << 32) | ((unsigned long long)(gridDim.x * gridDim.y * gridDim.z & 0x1FFFF) << 15) |
                              ((unsigned long long)(blockDim.x & 0x7FF) << 4) | (unsigned long long)(ev & 15);
        }
    }
}
// same, recorded by `thread` of block (bx, 0, 0): phase marks of a CTA other than the first (the longest one of a causal attention grid)
__device__ __forceinline__ void ktrace_event_at(int ev, int thread, int bx) {
    unsigned long long * t = c_ktrace;
    if (t != nullptr && ((threadIdx.x ^ thread) | (blockIdx.x ^ bx) | blockIdx.y | blockIdx.z) == 0) {
        unsigned long long now, gid;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        asm volatile("mov.u64 %0, %%gridid;" : "=l"(gid));
        const unsigned long long slot = atomicAdd(t, 1ull);
        if (slot < KTRACE_CAP) {
            t[1 + 2 * slot] = now;
            t[2 + 2 * slot] = (gid << 32) | ((unsigned long long)(gridDim.x * gridDim.y * gridDim.z & 0x1FFFF) << 15) |
                              ((unsigned long long)(blockDim.x & 0x7FF) << 4) | (unsigned long long)(ev & 15);
        }
    }
}
void ktrace_register(int (*setter)(unsigned long long *));
namespace {
struct ktrace_registrar {
    ktrace_registrar() {
        ktrace_register([](unsigned long long * p) -> int { return (int)cudaMemcpyToSymbol(c_ktrace, &p, sizeof(p)); });
    }
};
static ktrace_registrar ktrace_registrar_instance;
}
#endif

// programmatic dependent launch: wait for the producer grid / allow the dependent grid to start
__device__ __forceinline__ void pdl_wait()    { asm volatile("griddepcontrol.wait;" ::: "memory"); ktrace_event(1); }
__device__ __forceinline__ void pdl_trigger() { ktrace_event(0); asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// rms_norm arithmetic shared by rms_norm_kernel (glue.cu), add_rms_norm_mul_kernel (fused.cu) and the GEMV's fused prologue
// (mmvq_r.cuh): same operations in the same order, bit-identical results
// (explicit fused multiply-adds: left to the compiler, the same expression contracted as x*x first in one kernel and y*y
// first in the other -- a last-bit difference in one layer of one decode step out of hundreds, found by hashing node outputs)
__device__ __forceinline__ float sumsq4(const float4 v) {
    return __fmaf_rn(v.w, v.w, __fmaf_rn(v.z, v.z, __fmaf_rn(v.y, v.y, __fmul_rn(v.x, v.x))));
}
__device__ __forceinline__ float rms_scale(float tot, int K, float eps) { return rsqrtf(tot / (float)K + eps); }

// ---- row-split hand-over folded into a kernel (protocol of csrc/peer.cu; b200_mmvq_args.hs_*) ---------------------------
// A launch on a peer device waits for "the activations have arrived" in the prologue of every CTA and, from the CTA that
// finishes last, lowers that flag and raises "my slice of dst is written" in the main device's memory; the main device's own
// launch waits for those flags in its last CTA, so that the completion of its kernel means every slice is there.  Replaces
// three dependent launches per hand-over (wait kernel -> GEMV -> signal kernel) of the first version.
struct hs_params {
    const uint32_t * wait;               // every CTA: spin until != 0 before the activations are read
    uint32_t * clear;                    // last CTA: store 0 (the wait flag, once nobody reads it any more)
    uint32_t * signal;                   // last CTA: system fence, then store 1 (remote)
    uint32_t * join; int join_stride; uint32_t join_mask;      // last CTA: wait for join[t * stride], t in mask; lower them
    unsigned * counter;                  // ticket that elects the last CTA; returns to 0 by itself
    unsigned long long timeout_ns;       // spin budget (0: none); expiry sets *error and goes on (results are then garbage)
    uint32_t * error;
};
__device__ __forceinline__ uint32_t hs_ld_acquire_sys(const uint32_t * p) {
    uint32_t v; asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory"); return v;
}
__device__ __forceinline__ void hs_st_release_sys(uint32_t * p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" :: "l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void hs_spin(const uint32_t * f, const hs_params & h) {
    unsigned long long t0 = 0;
    if (h.timeout_ns) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    int n = 0;
    while (hs_ld_acquire_sys(f) == 0u) {
        if (++n > 64) __nanosleep(32);
        if (h.timeout_ns && (n & 255) == 0) {
            unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
            if (t - t0 > h.timeout_ns) { if (h.error) *(volatile uint32_t *)h.error = 1u; break; }
        }
    }
}
// one thread per CTA, before the activations are fetched (generic loads or a bulk copy issued by the same thread)
__device__ __forceinline__ void hs_wait_acts(const hs_params & h) {
    if (!h.wait) return;
    hs_spin(h.wait, h);
    asm volatile("fence.proxy.async;" ::: "memory");        // the bulk copy that follows reads what the flag guards
}
// one thread per CTA, after the threads of the CTA that stored results have synchronised with it (__syncwarp / bar.sync);
// total = CTAs of the launch.  Fences, measured between two B200s (tools/p2p_pingpong.cu, profiles/r02_p2p_pingpong.log): a
// system-scope fence after remote stores costs 1.9 us and 296 CTAs issuing one each serialise (29 us per hand-over); a gpu-scope
// fence costs 0.5 us.  So every CTA orders its stores before its ticket at gpu scope, and only the CTA that draws the last
// ticket -- whose atomic has observed all the others -- pays for the system-scope fence in front of the flag (9 us per hand-over
// with 296 CTAs, no stale reads): the fence is cumulative over the writes that happen-before it.
__device__ __forceinline__ void hs_finish(const hs_params & h, unsigned total) {
    if (!h.counter) return;
    __threadfence();
    if (atomicInc(h.counter, total - 1) != total - 1) return;
    for (uint32_t m = h.join_mask, t = 0; m; m >>= 1, ++t) {
        if (!(m & 1u)) continue;
        uint32_t * f = h.join + (size_t)t * h.join_stride;
        hs_spin(f, h);
        *(volatile uint32_t *)f = 0u;
    }
    if (h.clear) *(volatile uint32_t *)h.clear = 0u;
    __threadfence_system();
    if (h.signal) asm volatile("st.relaxed.sys.global.u32 [%0], %1;" :: "l"(h.signal), "r"(1u) : "memory");
}

extern int g_launches;   // kernels launched by this library (bench.py gpu_launches)
#define B200_HS_FROM_ARGS(h, a) do { (h).wait = (a)->hs_wait; (h).clear = (a)->hs_clear; (h).signal = (a)->hs_signal; (h).join = (a)->hs_join; \
    (h).join_stride = (a)->hs_join_stride; (h).join_mask = (a)->hs_join ? (a)->hs_join_mask : 0u; (h).counter = (a)->hs_counter; \
    (h).timeout_ns = (unsigned long long)(a)->hs_timeout_ms * 1000000ull; (h).error = (a)->hs_error; } while (0)
int pdl_default();       // env B200_PDL (default 1): launch kernels with programmatic dependent launch

}  // namespace b200

#define B200_CUDA_CHECK(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) return (int)e_; } while (0)
