// glue.cu -- the non-matmul nodes of the Llama / Mixtral decode and prefill graphs for B200.
// These are small HBM/latency-bound element-wise and row-wise kernels; every one takes part in the
// programmatic-dependent-launch chain (griddepcontrol.launch_dependents at entry, .wait before the
// first dependent access) so that the next GEMV can prefetch its weights while they run.
// Semantics follow the reference CPU operators (ggml/src/ggml-cpu/ops.cpp), cited per kernel.
#include "common.cuh"
#include "rope_units.cuh"
#include "../../include/b200_ops.h"

#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cuda_bf16.h>

namespace b200 {

struct tens {           // kernel-side copy of b200_tensor
    char * d;
    int64_t ne[4];
    int64_t nb[4];
};
static tens mk(const b200_tensor * t) {
    tens r; r.d = (char *)t->data;
    for (int i = 0; i < 4; ++i) { r.ne[i] = t->ne[i]; r.nb[i] = t->nb[i]; }
    return r;
}
static int64_t nelem(const b200_tensor * t) { return t->ne[0] * t->ne[1] * t->ne[2] * t->ne[3]; }
static int64_t nrows(const b200_tensor * t) { return t->ne[1] * t->ne[2] * t->ne[3]; }
static int type_size(int t) { return block_bytes(t); }
static bool is_contig(const b200_tensor * t) {
    int64_t s = type_size(t->type);
    if (block_elems(t->type) != 1) {
        if (t->nb[0] != s) return false;
        s = t->ne[0] / block_elems(t->type) * s;
        for (int i = 1; i < 4; ++i) { if (t->ne[i] != 1 && t->nb[i] != s) return false; s *= t->ne[i]; }
        return true;
    }
    for (int i = 0; i < 4; ++i) { if (t->ne[i] != 1 && t->nb[i] != s) return false; s *= t->ne[i]; }
    return true;
}
static bool same_shape(const b200_tensor * a, const b200_tensor * b) {
    for (int i = 0; i < 4; ++i) if (a->ne[i] != b->ne[i]) return false;
    return true;
}

// launch helper: PDL attribute on every glue kernel
template <typename... KArgs, typename... Args>
static int launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t lc = {};
    lc.gridDim = grid; lc.blockDim = block; lc.dynamicSmemBytes = smem; lc.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = pdl_default() ? 1 : 0;
    B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, kern, KArgs(args)...));
    ++g_launches;
    return 0;
}

__device__ __forceinline__ void pdl_enter() { pdl_trigger(); pdl_wait(); }

// ---- binary ops with broadcast (reference ggml-cpu/binary-ops.cpp; GPU analogue binbcast.cu) ----
template <int OP> __device__ __forceinline__ float binop(float a, float b) {
    if constexpr (OP == B200_ADD) return a + b;
    else if constexpr (OP == B200_SUB) return a - b;
    else if constexpr (OP == B200_MUL) return a * b;
    else return a / b;
}
template <int OP>
__global__ void __launch_bounds__(256) binary_kernel(const tens a, const tens b, const tens dst, int64_t n) {
    pdl_enter();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i0 = i % dst.ne[0], r = i / dst.ne[0];
        const int64_t i1 = r % dst.ne[1], r2 = r / dst.ne[1];
        const int64_t i2 = r2 % dst.ne[2], i3 = r2 / dst.ne[2];
        const float x = *(const float *)(a.d + i0 * a.nb[0] + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
        const float y = *(const float *)(b.d + (i0 % b.ne[0]) * b.nb[0] + (i1 % b.ne[1]) * b.nb[1] + (i2 % b.ne[2]) * b.nb[2] + (i3 % b.ne[3]) * b.nb[3]);
        *(float *)(dst.d + i0 * dst.nb[0] + i1 * dst.nb[1] + i2 * dst.nb[2] + i3 * dst.nb[3]) = binop<OP>(x, y);
    }
}
// fast path: everything contiguous, b either same shape or one row broadcast over rows; float4
template <int OP>
__global__ void __launch_bounds__(256) binary_rows_kernel(const float * __restrict__ a, const float * __restrict__ b, float * __restrict__ dst,
                                                          int64_t n4, int64_t brow4) {
    pdl_enter();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        const float4 x = reinterpret_cast<const float4 *>(a)[i];
        const float4 y = reinterpret_cast<const float4 *>(b)[i % brow4];
        float4 o; o.x = binop<OP>(x.x, y.x); o.y = binop<OP>(x.y, y.y); o.z = binop<OP>(x.z, y.z); o.w = binop<OP>(x.w, y.w);
        reinterpret_cast<float4 *>(dst)[i] = o;
    }
}

// ---- unary (reference ggml-cpu/unary-ops.cpp, vec.h) ----
template <int OP> __device__ __forceinline__ float unop(float x) {
    if constexpr (OP == B200_UNARY_ABS) return fabsf(x);
    else if constexpr (OP == B200_UNARY_SGN) return x > 0.f ? 1.f : (x < 0.f ? -1.f : 0.f);
    else if constexpr (OP == B200_UNARY_NEG) return -x;
    else if constexpr (OP == B200_UNARY_STEP) return x > 0.f ? 1.f : 0.f;
    else if constexpr (OP == B200_UNARY_TANH) return tanhf(x);
    else if constexpr (OP == B200_UNARY_ELU) return x > 0.f ? x : expm1f(x);
    else if constexpr (OP == B200_UNARY_RELU) return fmaxf(x, 0.f);
    else if constexpr (OP == B200_UNARY_SIGMOID) return 1.0f / (1.0f + expf(-x));
    else if constexpr (OP == B200_UNARY_GELU) { const float c = 0.044715f, s = 0.79788456080286535587989211986876f; return 0.5f * x * (1.0f + tanhf(s * x * (1.0f + c * x * x))); }
    else if constexpr (OP == B200_UNARY_GELU_QUICK) return x * (1.0f / (1.0f + expf(-1.702f * x)));
    else if constexpr (OP == B200_UNARY_SILU) return x / (1.0f + expf(-x));
    else if constexpr (OP == B200_UNARY_HARDSWISH) return x * fminf(1.0f, fmaxf(0.0f, (x + 3.0f) / 6.0f));
    else if constexpr (OP == B200_UNARY_HARDSIGMOID) return fminf(1.0f, fmaxf(0.0f, (x + 3.0f) / 6.0f));
    else if constexpr (OP == B200_UNARY_EXP) return expf(x);
    else return 0.5f * x * (1.0f + erff(x * 0.70710678118654752440f));
}
template <int OP>
__global__ void __launch_bounds__(256) unary_kernel(const float * __restrict__ x, float * __restrict__ y, int64_t n) {
    pdl_enter();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) y[i] = unop<OP>(x[i]);
}
__global__ void __launch_bounds__(256) scale_kernel(const float * __restrict__ x, float * __restrict__ y, float s, int64_t n) {
    pdl_enter();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) y[i] = x[i] * s;
}

// ---- rms_norm (reference ggml-cpu/ops.cpp:3285-3312): one CTA per row ----
// The row is read once with independent 128-bit loads (all in flight together: a decode step has a
// single row, so the kernel is pure load latency) and kept in registers for the scaling pass.
template <int VPT>      // float4 per thread
__global__ void __launch_bounds__(1024) rms_norm_kernel(const tens src, const tens dst, float eps) {
    pdl_enter();
    const int64_t r = blockIdx.x;
    const int64_t i1 = r % src.ne[1], i2 = (r / src.ne[1]) % src.ne[2], i3 = r / (src.ne[1] * src.ne[2]);
    const float4 * x = (const float4 *)(src.d + i1 * src.nb[1] + i2 * src.nb[2] + i3 * src.nb[3]);
    float4 * y = (float4 *)(dst.d + i1 * dst.nb[1] + i2 * dst.nb[2] + i3 * dst.nb[3]);
    const int n4 = (int)(src.ne[0] >> 2);
    float4 v[VPT];
#pragma unroll
    for (int k = 0; k < VPT; ++k) { const int i = threadIdx.x + k * blockDim.x; v[k] = i < n4 ? x[i] : make_float4(0.f, 0.f, 0.f, 0.f); }
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < VPT; ++k) s += sumsq4(v[k]);
    s = warp_sum(s);
    __shared__ float part[32];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
    __syncthreads();
    float tot = 0.0f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) tot += part[w];
    const float scale = rms_scale(tot, (int)src.ne[0], eps);
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
        const int i = threadIdx.x + k * blockDim.x;
        if (i < n4) y[i] = make_float4(v[k].x * scale, v[k].y * scale, v[k].z * scale, v[k].w * scale);
    }
}
// generic fallback (unaligned rows or very long rows)
__global__ void __launch_bounds__(256) rms_norm_generic_kernel(const tens src, const tens dst, float eps) {
    pdl_enter();
    const int64_t r = blockIdx.x;
    const int64_t i1 = r % src.ne[1], i2 = (r / src.ne[1]) % src.ne[2], i3 = r / (src.ne[1] * src.ne[2]);
    const float * x = (const float *)(src.d + i1 * src.nb[1] + i2 * src.nb[2] + i3 * src.nb[3]);
    float * y = (float *)(dst.d + i1 * dst.nb[1] + i2 * dst.nb[2] + i3 * dst.nb[3]);
    const int n = (int)src.ne[0];
    float s = 0.0f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) { const float v = x[i]; s += v * v; }
    s = warp_sum(s);
    __shared__ float part[8];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
    __syncthreads();
    float tot = 0.0f;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) tot += part[w];
    const float scale = rsqrtf(tot / n + eps);
    for (int i = threadIdx.x; i < n; i += blockDim.x) y[i] = x[i] * scale;
}

// ---- rope (reference ggml-cpu/ops.cpp:5080-5118, 5178-5360): modes normal and neox ----
template <typename T>
__global__ void __launch_bounds__(256) rope_kernel(const tens src, const tens dst, const int32_t * __restrict__ pos,
                                                   const float * __restrict__ ff, const rope_dev p) {
    pdl_enter();
    const int64_t half0 = src.ne[0] / 2;
    const int64_t total = half0 * src.ne[1] * src.ne[2] * src.ne[3];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t ip = i % half0, r = i / half0;
        const int64_t i1 = r % src.ne[1], i2 = (r / src.ne[1]) % src.ne[2], i3 = r / (src.ne[1] * src.ne[2]);
        const char * s = src.d + i1 * src.nb[1] + i2 * src.nb[2] + i3 * src.nb[3];
        char * d = dst.d + i1 * dst.nb[1] + i2 * dst.nb[2] + i3 * dst.nb[3];
        const int64_t i0 = 2 * ip;
        if (i0 >= p.n_dims) {   // pass-through of the unrotated tail
            ((T *)d)[i0] = ((const T *)s)[i0]; ((T *)d)[i0 + 1] = ((const T *)s)[i0 + 1];
            continue;
        }
        float c, sn; rope_cs(p, ip, pos[i2], ff, c, sn);
        const int64_t ia = p.neox ? ip : i0, ib = p.neox ? ip + p.n_dims / 2 : i0 + 1;
        const float x0 = (float)((const T *)s)[ia], x1 = (float)((const T *)s)[ib];
        float ya, yb; rope_pair(x0, x1, c, sn, ya, yb);
        ((T *)d)[ia] = (T)ya;
        ((T *)d)[ib] = (T)yb;
    }
}

// Prefill shape (many tokens): one thread per (token, pair, group of ROPE_HPT heads), so the powf / sincos of rope_cs are paid once
// per group instead of once per head, and the group's loads are all issued before the first use (the per-element kernel is bound
// by the transcendentals, a thread that walks all the heads by load latency).  Same rope_cs / rope_pair arithmetic: bit-identical.
constexpr int ROPE_HPT = 8;
template <typename T>
__global__ void __launch_bounds__(256) rope_heads_kernel(const tens src, const tens dst, const int32_t * __restrict__ pos,
                                                         const float * __restrict__ ff, const rope_dev p, int ngrp) {
    pdl_enter();
    const int64_t half0 = src.ne[0] / 2;
    const int64_t total = half0 * ngrp * src.ne[2] * src.ne[3];
    const int nh = (int)src.ne[1];
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t ip = i % half0, r = i / half0;
        const int h0 = (int)(r % ngrp) * ROPE_HPT;
        const int64_t r2 = r / ngrp, i2 = r2 % src.ne[2], i3 = r2 / src.ne[2];
        const char * s = src.d + i2 * src.nb[2] + i3 * src.nb[3] + h0 * src.nb[1];
        char * d = dst.d + i2 * dst.nb[2] + i3 * dst.nb[3] + h0 * dst.nb[1];
        const int64_t i0 = 2 * ip;
        const bool tail = i0 >= p.n_dims;          // pass-through of the unrotated tail
        const int64_t ia = tail ? i0 : (p.neox ? ip : i0), ib = tail ? i0 + 1 : (p.neox ? ip + p.n_dims / 2 : i0 + 1);
        float x0[ROPE_HPT], x1[ROPE_HPT];
#pragma unroll
        for (int h = 0; h < ROPE_HPT; ++h)
            if (h0 + h < nh) {
                const T * sh = (const T *)(s + h * src.nb[1]);
                x0[h] = (float)sh[ia]; x1[h] = (float)sh[ib];
            }
        float c = 1.0f, sn = 0.0f;
        if (!tail) rope_cs(p, ip, pos[i2], ff, c, sn);
#pragma unroll
        for (int h = 0; h < ROPE_HPT; ++h)
            if (h0 + h < nh) {
                T * dh = (T *)(d + h * dst.nb[1]);
                float ya = x0[h], yb = x1[h];
                if (!tail) rope_pair(x0[h], x1[h], c, sn, ya, yb);
                dh[ia] = (T)ya;
                dh[ib] = (T)yb;
            }
    }
}

// ---- cpy / cont / dup with conversion ----
template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// element i (flat index in the logical order of its own shape) -> byte offset
__device__ __forceinline__ int64_t flat_off(const tens & t, int64_t i) {
    const int64_t i0 = i % t.ne[0], r = i / t.ne[0];
    const int64_t i1 = r % t.ne[1], r2 = r / t.ne[1];
    return i0 * t.nb[0] + i1 * t.nb[1] + (r2 % t.ne[2]) * t.nb[2] + (r2 / t.ne[2]) * t.nb[3];
}
template <typename S, typename D>
__global__ void __launch_bounds__(256) cpy_kernel(const tens src, const tens dst, int64_t n, char * const * dslot) {
    pdl_enter();
    char * dd = dslot ? *dslot : dst.d;          // CUDA-graph replay: the destination (KV cache cell) comes from a pointer table
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x)
        *(D *)(dd + flat_off(dst, i)) = from_f32<D>(to_f32<S>(*(const S *)(src.d + flat_off(src, i))));
}
// four consecutive elements per thread when rows of both sides hold a multiple of four and every row start is vector aligned
// (the K / V store of a prefill batch): one offset computation per four elements, 16-byte loads, 8- or 16-byte stores
template <typename S, typename D>
__global__ void __launch_bounds__(256) cpy_vec4_kernel(const tens src, const tens dst, int64_t n4) {
    pdl_enter();
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (int64_t)gridDim.x * blockDim.x) {
        struct alignas(4 * sizeof(S)) sv_t { S v[4]; } sv = *(const sv_t *)(src.d + flat_off(src, 4 * i));
        struct alignas(4 * sizeof(D)) dv_t { D v[4]; } dv;
#pragma unroll
        for (int j = 0; j < 4; ++j) dv.v[j] = from_f32<D>(to_f32<S>(sv.v[j]));
        *(dv_t *)(dst.d + flat_off(dst, 4 * i)) = dv;
    }
}
// f32 -> q8_0 blocks: one warp per block of 32 consecutive source elements (the same flat order);
// arithmetic of quantize_row_q8_0_ref (ggml-quants.c:191-212): d = amax/127, id = 1/d, q = roundf(x*id)
__global__ void __launch_bounds__(256) cpy_f32_q8_0_kernel(const tens src, const tens dst, int64_t nblocks, char * const * dslot) {
    pdl_enter();
    char * dd = dslot ? *dslot : dst.d;
    const int lane = threadIdx.x & 31;
    for (int64_t b = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; b < nblocks; b += ((int64_t)gridDim.x * blockDim.x) >> 5) {
        const float x = *(const float *)(src.d + flat_off(src, b * 32 + lane));
        const float amax = warp_max(fabsf(x));
        const float d = __fdiv_rn(amax, 127.0f);
        const float id = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
        const int q = (int)roundf(x * id);
        // destination block: dst is laid out in blocks along ne[0]
        const int64_t bpr = dst.ne[0] / 32, i0 = b % bpr, r = b / bpr;
        const int64_t i1 = r % dst.ne[1], r2 = r / dst.ne[1];
        uint8_t * blk = (uint8_t *)(dd + i0 * dst.nb[0] + i1 * dst.nb[1] + (r2 % dst.ne[2]) * dst.nb[2] + (r2 / dst.ne[2]) * dst.nb[3]);
        blk[2 + lane] = (uint8_t)(int8_t)q;
        if (lane == 0) *reinterpret_cast<__half *>(blk) = __float2half_rn(d);
    }
}
// same-type copy of block data (quantised or not) between strided tensors of identical shape
__global__ void __launch_bounds__(256) cpy_blocks_kernel(const tens src, const tens dst, int64_t nblocks, int bb, int be, char * const * dslot) {
    pdl_enter();
    char * dd = dslot ? *dslot : dst.d;
    for (int64_t b = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; b < nblocks; b += (int64_t)gridDim.x * blockDim.x) {
        const int64_t bps = src.ne[0] / be, s0 = b % bps, sr = b / bps;
        const int64_t s1 = sr % src.ne[1], sr2 = sr / src.ne[1];
        const char * sp = src.d + s0 * src.nb[0] + s1 * src.nb[1] + (sr2 % src.ne[2]) * src.nb[2] + (sr2 / src.ne[2]) * src.nb[3];
        const int64_t bpd = dst.ne[0] / be, d0 = b % bpd, dr = b / bpd;
        const int64_t d1 = dr % dst.ne[1], dr2 = dr / dst.ne[1];
        char * dp = dd + d0 * dst.nb[0] + d1 * dst.nb[1] + (dr2 % dst.ne[2]) * dst.nb[2] + (dr2 / dst.ne[2]) * dst.nb[3];
        for (int k = 0; k < bb; k += 2) *(uint16_t *)(dp + k) = *(const uint16_t *)(sp + k);
    }
}

// ---- several rope / KV-store nodes of a decode step in one launch (blockIdx.y = task) ----
// kind 0: dst = rope(src);  kind 1: dst = rope(src) and dst2 = convert(dst) in cpy's flat element order;  kind 2: dst2 = convert(src)
struct ew_task { int kind, d2type; tens src, dst, dst2; const int32_t * pos; const float * ff; rope_dev rp; char * const * dslot; int64_t n;
                 int vec4, ngrp; int64_t nwide; };      // prefill form: 4-wide copy possible, head groups per (token, pair), threads needed
struct ew_batch { ew_task t[4]; };
__device__ __forceinline__ void store_conv(char * base, const tens & d2, int d2type, int64_t flat, float v) {
    char * p = base + flat_off(d2, flat);
    if (d2type == T_F16) *(__half *)p = __float2half_rn(v);
    else if (d2type == T_BF16) *(__nv_bfloat16 *)p = __float2bfloat16_rn(v);
    else *(float *)p = v;
}
__global__ void __launch_bounds__(256) ew_batch_kernel(const __grid_constant__ ew_batch b) {
    pdl_enter();
    const ew_task & t = b.t[blockIdx.y];
    char * d2 = t.kind ? (t.dslot ? *t.dslot : t.dst2.d) : nullptr;
    if (t.kind == 2) {
        for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < t.n; i += (int64_t)gridDim.x * blockDim.x)
            store_conv(d2, t.dst2, t.d2type, i, *(const float *)(t.src.d + flat_off(t.src, i)));
        return;
    }
    const tens & src = t.src; const tens & dst = t.dst; const rope_dev & p = t.rp;
    const int64_t half0 = src.ne[0] / 2;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < t.n; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t ip = i % half0, r = i / half0;
        const int64_t i1 = r % src.ne[1], i2 = (r / src.ne[1]) % src.ne[2], i3 = r / (src.ne[1] * src.ne[2]);
        const float * s = (const float *)(src.d + i1 * src.nb[1] + i2 * src.nb[2] + i3 * src.nb[3]);
        float * d = (float *)(dst.d + i1 * dst.nb[1] + i2 * dst.nb[2] + i3 * dst.nb[3]);
        const int64_t row0 = r * src.ne[0];          // flat index of the row's first element (cpy order)
        const int64_t i0 = 2 * ip;
        int64_t ia, ib; float ya, yb;
        if (i0 >= p.n_dims) { ia = i0; ib = i0 + 1; ya = s[ia]; yb = s[ib]; }
        else {
            float c, sn; rope_cs(p, ip, t.pos[i2], t.ff, c, sn);
            ia = p.neox ? ip : i0; ib = p.neox ? ip + p.n_dims / 2 : i0 + 1;
            const float x0 = s[ia], x1 = s[ib];
            rope_pair(x0, x1, c, sn, ya, yb);
        }
        d[ia] = ya; d[ib] = yb;
        if (t.kind == 1) { store_conv(d2, t.dst2, t.d2type, row0 + ia, ya); store_conv(d2, t.dst2, t.d2type, row0 + ib, yb); }
    }
}

// The same tasks for a prefill batch (thousands of tokens): rope walks ROPE_HPT heads per thread (rope_heads_kernel's mapping: cos/sin
// once per head group), the plain store moves four elements per thread (cpy_vec4_kernel's) -- one launch for rope(q), rope(k) -> K
// cache and v -> V cache instead of four.  Same arithmetic and roundings as the kernels it stands for.
__global__ void __launch_bounds__(256) ew_batch_wide_kernel(const __grid_constant__ ew_batch b) {
    pdl_enter();
    const ew_task & t = b.t[blockIdx.y];
    char * d2 = t.kind ? (t.dslot ? *t.dslot : t.dst2.d) : nullptr;
    if (t.kind == 2) {
        if (t.vec4) {
            for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < t.nwide; i += (int64_t)gridDim.x * blockDim.x) {
                const float4 v = *(const float4 *)(t.src.d + flat_off(t.src, 4 * i));
                char * o = d2 + flat_off(t.dst2, 4 * i);
                if (t.d2type == T_F16) { __half2 a = __floats2half2_rn(v.x, v.y), c = __floats2half2_rn(v.z, v.w); uint2 w; w.x = *(uint32_t *)&a; w.y = *(uint32_t *)&c; *(uint2 *)o = w; }
                else if (t.d2type == T_BF16) { __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), c = __floats2bfloat162_rn(v.z, v.w); uint2 w; w.x = *(uint32_t *)&a; w.y = *(uint32_t *)&c; *(uint2 *)o = w; }
                else *(float4 *)o = v;
            }
        } else {
            for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < t.n; i += (int64_t)gridDim.x * blockDim.x)
                store_conv(d2, t.dst2, t.d2type, i, *(const float *)(t.src.d + flat_off(t.src, i)));
        }
        return;
    }
    const tens & src = t.src; const tens & dst = t.dst; const rope_dev & p = t.rp;
    const int64_t half0 = src.ne[0] / 2;
    const int nh = (int)src.ne[1], ngrp = t.ngrp;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < t.nwide; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t ip = i % half0, r = i / half0;
        const int h0 = (int)(r % ngrp) * ROPE_HPT;
        const int64_t r2 = r / ngrp, i2 = r2 % src.ne[2], i3 = r2 / src.ne[2];
        const char * sb = src.d + i2 * src.nb[2] + i3 * src.nb[3] + h0 * src.nb[1];
        char * db = dst.d + i2 * dst.nb[2] + i3 * dst.nb[3] + h0 * dst.nb[1];
        const int64_t row0 = ((i3 * src.ne[2] + i2) * nh + h0) * src.ne[0];      // flat index of head h0's first element (cpy order)
        const int64_t i0 = 2 * ip;
        const bool tail = i0 >= p.n_dims;
        const int64_t ia = tail ? i0 : (p.neox ? ip : i0), ib = tail ? i0 + 1 : (p.neox ? ip + p.n_dims / 2 : i0 + 1);
        float x0[ROPE_HPT], x1[ROPE_HPT];
#pragma unroll
        for (int h = 0; h < ROPE_HPT; ++h)
            if (h0 + h < nh) {
                const float * sh = (const float *)(sb + h * src.nb[1]);
                x0[h] = sh[ia]; x1[h] = sh[ib];
            }
        float c = 1.0f, sn = 0.0f;
        if (!tail) rope_cs(p, ip, t.pos[i2], t.ff, c, sn);
#pragma unroll
        for (int h = 0; h < ROPE_HPT; ++h)
            if (h0 + h < nh) {
                float * dh = (float *)(db + h * dst.nb[1]);
                float ya = x0[h], yb = x1[h];
                if (!tail) rope_pair(x0[h], x1[h], c, sn, ya, yb);
                dh[ia] = ya; dh[ib] = yb;
                if (t.kind == 1) {
                    store_conv(d2, t.dst2, t.d2type, row0 + h * src.ne[0] + ia, ya);
                    store_conv(d2, t.dst2, t.d2type, row0 + h * src.ne[0] + ib, yb);
                }
            }
    }
}

static unsigned grid_for(int64_t n, int per_block) {
    return (unsigned)std::max<int64_t>(1, std::min<int64_t>((n + per_block - 1) / per_block, 148 * 8));
}

}  // namespace b200

using namespace b200;

extern "C" int b200_op_binary(int op, const b200_tensor * a, const b200_tensor * b, const b200_tensor * dst, void * stream) {
    if (a->type != T_F32 || b->type != T_F32 || dst->type != T_F32 || !same_shape(a, dst)) return -1;
    for (int i = 0; i < 4; ++i) if (b->ne[i] <= 0 || dst->ne[i] % b->ne[i]) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t n = nelem(dst);
    if (n == 0) return 0;
    const int64_t brow = nelem(b);
    const bool fast = is_contig(a) && is_contig(b) && is_contig(dst) && n % 4 == 0 && brow % 4 == 0 &&
                      ((((uintptr_t)a->data | (uintptr_t)b->data | (uintptr_t)dst->data) & 15) == 0) &&
                      (same_shape(a, b) || (b->ne[0] == a->ne[0] && brow == b->ne[0]));
    if (fast) {
        const unsigned g = grid_for(n / 4, 256);
        switch (op) {
            case B200_ADD: return launch(binary_rows_kernel<B200_ADD>, g, 256, 0, st, (const float *)a->data, (const float *)b->data, (float *)dst->data, n / 4, brow / 4);
            case B200_SUB: return launch(binary_rows_kernel<B200_SUB>, g, 256, 0, st, (const float *)a->data, (const float *)b->data, (float *)dst->data, n / 4, brow / 4);
            case B200_MUL: return launch(binary_rows_kernel<B200_MUL>, g, 256, 0, st, (const float *)a->data, (const float *)b->data, (float *)dst->data, n / 4, brow / 4);
            case B200_DIV: return launch(binary_rows_kernel<B200_DIV>, g, 256, 0, st, (const float *)a->data, (const float *)b->data, (float *)dst->data, n / 4, brow / 4);
        }
        return -1;
    }
    const unsigned g = grid_for(n, 256);
    switch (op) {
        case B200_ADD: return launch(binary_kernel<B200_ADD>, g, 256, 0, st, mk(a), mk(b), mk(dst), n);
        case B200_SUB: return launch(binary_kernel<B200_SUB>, g, 256, 0, st, mk(a), mk(b), mk(dst), n);
        case B200_MUL: return launch(binary_kernel<B200_MUL>, g, 256, 0, st, mk(a), mk(b), mk(dst), n);
        case B200_DIV: return launch(binary_kernel<B200_DIV>, g, 256, 0, st, mk(a), mk(b), mk(dst), n);
    }
    return -1;
}

#define B200_UNARY_CASE(OP) case OP: return launch(unary_kernel<OP>, g, 256, 0, st, (const float *)src->data, (float *)dst->data, n);
extern "C" int b200_op_unary(int op, const b200_tensor * src, const b200_tensor * dst, void * stream) {
    if (src->type != T_F32 || dst->type != T_F32 || !same_shape(src, dst) || !is_contig(src) || !is_contig(dst)) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t n = nelem(dst);
    if (n == 0) return 0;
    const unsigned g = grid_for(n, 256);
    switch (op) {
        B200_UNARY_CASE(B200_UNARY_ABS) B200_UNARY_CASE(B200_UNARY_SGN) B200_UNARY_CASE(B200_UNARY_NEG) B200_UNARY_CASE(B200_UNARY_STEP)
        B200_UNARY_CASE(B200_UNARY_TANH) B200_UNARY_CASE(B200_UNARY_ELU) B200_UNARY_CASE(B200_UNARY_RELU) B200_UNARY_CASE(B200_UNARY_SIGMOID)
        B200_UNARY_CASE(B200_UNARY_GELU) B200_UNARY_CASE(B200_UNARY_GELU_QUICK) B200_UNARY_CASE(B200_UNARY_SILU)
        B200_UNARY_CASE(B200_UNARY_HARDSWISH) B200_UNARY_CASE(B200_UNARY_HARDSIGMOID) B200_UNARY_CASE(B200_UNARY_EXP)
        B200_UNARY_CASE(B200_UNARY_GELU_ERF)
    }
    return -1;
}

extern "C" int b200_op_scale(const b200_tensor * src, const b200_tensor * dst, float scale, void * stream) {
    if (src->type != T_F32 || dst->type != T_F32 || !same_shape(src, dst) || !is_contig(src) || !is_contig(dst)) return -1;
    const int64_t n = nelem(dst);
    if (n == 0) return 0;
    return launch(scale_kernel, grid_for(n, 256), 256, 0, (cudaStream_t)stream, (const float *)src->data, (float *)dst->data, scale, n);
}

extern "C" int b200_op_rms_norm(const b200_tensor * src, const b200_tensor * dst, float eps, void * stream) {
    if (src->type != T_F32 || dst->type != T_F32 || !same_shape(src, dst) || src->nb[0] != 4 || dst->nb[0] != 4) return -1;
    const int64_t r = nrows(src);
    if (r == 0 || src->ne[0] == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    const bool al = src->ne[0] % 4 == 0 && ((((uintptr_t)src->data | (uintptr_t)dst->data | src->nb[1] | src->nb[2] | src->nb[3] |
                                              dst->nb[1] | dst->nb[2] | dst->nb[3]) & 15) == 0);
    const int64_t n4 = src->ne[0] / 4;
    if (al && n4 <= 4096) {
        // few rows (decode): many threads, one float4 each; many rows (prefill): 256 threads, up to 4 float4 each
        if (r <= 8 || n4 > 1024) {
            const int threads = (int)std::min<int64_t>(1024, (n4 + 31) / 32 * 32);
            if (n4 <= threads) return launch(rms_norm_kernel<1>, (unsigned)r, threads, 0, st, mk(src), mk(dst), eps);
            if (n4 <= 2 * threads) return launch(rms_norm_kernel<2>, (unsigned)r, threads, 0, st, mk(src), mk(dst), eps);
            return launch(rms_norm_kernel<4>, (unsigned)r, threads, 0, st, mk(src), mk(dst), eps);
        }
        const int threads = (int)std::min<int64_t>(256, (n4 + 31) / 32 * 32);
        if (n4 <= threads) return launch(rms_norm_kernel<1>, (unsigned)r, threads, 0, st, mk(src), mk(dst), eps);
        if (n4 <= 2 * threads) return launch(rms_norm_kernel<2>, (unsigned)r, threads, 0, st, mk(src), mk(dst), eps);
        return launch(rms_norm_kernel<4>, (unsigned)r, threads, 0, st, mk(src), mk(dst), eps);
    }
    return launch(rms_norm_generic_kernel, (unsigned)r, 256, 0, st, mk(src), mk(dst), eps);
}

extern "C" int b200_op_rope(const b200_tensor * src, const b200_tensor * pos, const b200_tensor * ffac, const b200_tensor * dst,
                            const b200_rope_params * prm, void * stream) {
    if (src->type != dst->type || (src->type != T_F32 && src->type != T_F16) || !same_shape(src, dst)) return -1;
    if (prm->mode != 0 && prm->mode != 2) return -1;               // normal and neox only (no mrope / vision)
    if (src->nb[0] != type_size(src->type) || dst->nb[0] != type_size(dst->type) || pos->type != T_I32) return -1;
    if (prm->n_dims % 2 || prm->n_dims > src->ne[0] || src->ne[0] % 2) return -1;
    const int64_t n = nelem(src) / 2;
    if (n == 0) return 0;
    rope_dev p;
    p.n_dims = prm->n_dims; p.neox = prm->mode == 2; p.freq_scale = prm->freq_scale; p.ext_factor = prm->ext_factor;
    p.attn_factor = prm->attn_factor; p.theta_scale = powf(prm->freq_base, -2.0f / prm->n_dims);
    // yarn correction dims (reference ggml.c:3770-3784)
    auto corr_dim = [&](float n_rot) { return prm->n_dims * logf(prm->n_ctx_orig / (n_rot * 2 * 3.14159265358979323846f)) / (2 * logf(prm->freq_base)); };
    p.corr0 = std::max(0.0f, floorf(corr_dim(prm->beta_fast)));
    p.corr1 = std::min((float)prm->n_dims - 1, ceilf(corr_dim(prm->beta_slow)));
    const float * ff = ffac ? (const float *)ffac->data : nullptr;
    // a prefill batch: enough (token, pair) positions to fill the GPU with one thread per ROPE_HPT heads
    const int ngrp = (int)((src->ne[1] + ROPE_HPT - 1) / ROPE_HPT);
    const int64_t npos = src->ne[0] / 2 * src->ne[2] * src->ne[3];
    static const bool heads_on = !getenv("B200_ROPE_HEADS") || atoi(getenv("B200_ROPE_HEADS")) != 0;
    if (heads_on && src->ne[1] >= 4 && npos >= 8192) {
        if (src->type == T_F32) return launch(rope_heads_kernel<float>, grid_for(npos * ngrp, 256), 256, 0, (cudaStream_t)stream, mk(src), mk(dst), (const int32_t *)pos->data, ff, p, ngrp);
        return launch(rope_heads_kernel<__half>, grid_for(npos * ngrp, 256), 256, 0, (cudaStream_t)stream, mk(src), mk(dst), (const int32_t *)pos->data, ff, p, ngrp);
    }
    if (src->type == T_F32) return launch(rope_kernel<float>, grid_for(n, 256), 256, 0, (cudaStream_t)stream, mk(src), mk(dst), (const int32_t *)pos->data, ff, p);
    return launch(rope_kernel<__half>, grid_for(n, 256), 256, 0, (cudaStream_t)stream, mk(src), mk(dst), (const int32_t *)pos->data, ff, p);
}

extern "C" int b200_op_cpy_supported(int s, int d) {
    const bool sf = s == T_F32 || s == T_F16 || s == B200_TYPE_BF16, df = d == T_F32 || d == T_F16 || d == B200_TYPE_BF16;
    if (sf && df) return 1;
    if (s == T_F32 && d == T_Q8_0) return 1;
    if (s == d && block_bytes(s) > 0) return 1;
    return 0;
}

#define B200_CPY_CASE(ST, SC, DT, DC) if (src->type == ST && dst->type == DT) return launch(cpy_kernel<SC, DC>, grid_for(n, 256), 256, 0, st, mk(src), mk(dst), n, dslot);
extern "C" int b200_op_cpy(const b200_tensor * src, const b200_tensor * dst, void * stream) { return b200_op_cpy_indirect(src, dst, nullptr, stream); }
extern "C" int b200_op_cpy_indirect(const b200_tensor * src, const b200_tensor * dst, void * const * dst_slot, void * stream) {
    if (!b200_op_cpy_supported(src->type, dst->type) || nelem(src) != nelem(dst)) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    char * const * dslot = (char * const *)dst_slot;
    const int64_t n = nelem(src);
    if (n == 0) return 0;
    if (!dslot && src->type == dst->type && is_contig(src) && is_contig(dst)) {
        const size_t bytes = (size_t)(n / block_elems(src->type)) * block_bytes(src->type);
        B200_CUDA_CHECK(cudaMemcpyAsync(dst->data, src->data, bytes, cudaMemcpyDeviceToDevice, st));
        return 0;
    }
    if (src->type == T_F32 && dst->type == T_Q8_0) {
        if (src->ne[0] % 32 || dst->ne[0] % 32) return -1;
        const int64_t nb = n / 32;
        return launch(cpy_f32_q8_0_kernel, grid_for(nb, 8), 256, 0, st, mk(src), mk(dst), nb, dslot);
    }
    if (src->type == dst->type && block_elems(src->type) > 1) {
        const int be = block_elems(src->type);
        if (src->ne[0] % be || dst->ne[0] % be) return -1;
        return launch(cpy_blocks_kernel, grid_for(n / be, 256), 256, 0, st, mk(src), mk(dst), n / be, block_bytes(src->type), be, dslot);
    }
    {   // vector path (see cpy_vec4_kernel)
        auto vec_ok = [](const b200_tensor * t) {
            const size_t a = 4 * type_size(t->type);
            return t->nb[0] == type_size(t->type) && t->ne[0] % 4 == 0 && ((uintptr_t)t->data % a) == 0 &&
                   (t->ne[1] == 1 || t->nb[1] % a == 0) && (t->ne[2] == 1 || t->nb[2] % a == 0) && (t->ne[3] == 1 || t->nb[3] % a == 0);
        };
        static const bool vec_on = !getenv("B200_CPY_VEC") || atoi(getenv("B200_CPY_VEC")) != 0;
        if (vec_on && !dslot && n >= 16384 && vec_ok(src) && vec_ok(dst)) {
#define B200_CPY_VCASE(ST, SC, DT, DC) if (src->type == ST && dst->type == DT) return launch(cpy_vec4_kernel<SC, DC>, grid_for(n / 4, 256), 256, 0, st, mk(src), mk(dst), n / 4);
            B200_CPY_VCASE(T_F32, float, T_F16, __half) B200_CPY_VCASE(T_F32, float, B200_TYPE_BF16, __nv_bfloat16)
            B200_CPY_VCASE(T_F16, __half, T_F32, float) B200_CPY_VCASE(T_F32, float, T_F32, float) B200_CPY_VCASE(T_F16, __half, T_F16, __half)
#undef B200_CPY_VCASE
        }
    }
    B200_CPY_CASE(T_F32, float, T_F32, float) B200_CPY_CASE(T_F32, float, T_F16, __half) B200_CPY_CASE(T_F32, float, B200_TYPE_BF16, __nv_bfloat16)
    B200_CPY_CASE(T_F16, __half, T_F32, float) B200_CPY_CASE(T_F16, __half, T_F16, __half) B200_CPY_CASE(T_F16, __half, B200_TYPE_BF16, __nv_bfloat16)
    B200_CPY_CASE(B200_TYPE_BF16, __nv_bfloat16, T_F32, float) B200_CPY_CASE(B200_TYPE_BF16, __nv_bfloat16, T_F16, __half)
    B200_CPY_CASE(B200_TYPE_BF16, __nv_bfloat16, B200_TYPE_BF16, __nv_bfloat16)
    return -1;
}

#define EW_FAIL() do { if (getenv("B200_DEBUG")) fprintf(stderr, "b200_op_rope_cpy_batch: refused at line %d\n", __LINE__); return -1; } while (0)
static int ew_build(const b200_ew_desc * tasks, int n, ew_batch & b, int64_t & nmax) {
    if (n < 1 || n > 4) EW_FAIL();
    nmax = 0;
    for (int i = 0; i < n; ++i) {
        const b200_ew_desc & d = tasks[i];
        ew_task & t = b.t[i];
        t.kind = d.kind;
        if (d.kind < 0 || d.kind > 2 || !d.src || d.src->type != T_F32 || (d.kind != 2 && d.src->nb[0] != 4)) EW_FAIL();
        t.src = mk(d.src);
        if (d.kind != 2) {
            const b200_rope_params * prm = d.rope;
            if (!d.dst || !prm || !d.pos || d.dst->type != T_F32 || d.dst->nb[0] != 4 || !same_shape(d.src, d.dst) || d.pos->type != T_I32) EW_FAIL();
            if ((prm->mode != 0 && prm->mode != 2) || prm->n_dims % 2 || prm->n_dims > d.src->ne[0] || d.src->ne[0] % 2) EW_FAIL();
            t.dst = mk(d.dst);
            t.pos = (const int32_t *)d.pos->data; t.ff = d.ffac ? (const float *)d.ffac->data : nullptr;
            t.rp = rope_dev_from(*prm);
            t.n = nelem(d.src) / 2;
            t.ngrp = (int)((d.src->ne[1] + ROPE_HPT - 1) / ROPE_HPT);
            t.nwide = d.src->ne[0] / 2 * t.ngrp * d.src->ne[2] * d.src->ne[3];
        } else { t.n = nelem(d.src); t.nwide = t.n / 4; }
        if (d.kind != 0) {
            if (!d.dst2 || nelem(d.dst2) != nelem(d.src)) EW_FAIL();
            if (d.dst2->type != T_F16 && d.dst2->type != T_F32 && d.dst2->type != T_BF16) EW_FAIL();
            if (d.dst2->nb[0] != type_size(d.dst2->type)) EW_FAIL();
            t.dst2 = mk(d.dst2); t.d2type = d.dst2->type; t.dslot = (char * const *)d.dst2_slot;
            if (d.kind == 2) {
                auto vec_ok = [](const b200_tensor * x) {
                    const size_t a = 4 * type_size(x->type);
                    return x->nb[0] == type_size(x->type) && x->ne[0] % 4 == 0 && ((uintptr_t)x->data % a) == 0 &&
                           (x->ne[1] == 1 || x->nb[1] % a == 0) && (x->ne[2] == 1 || x->nb[2] % a == 0) && (x->ne[3] == 1 || x->nb[3] % a == 0);
                };
                t.vec4 = !d.dst2_slot && vec_ok(d.src) && vec_ok(d.dst2);
            }
        }
        nmax = std::max(nmax, t.n);
    }
    return 0;
}
extern "C" int b200_op_rope_cpy_batch_check(const b200_ew_desc * tasks, int n) {
    ew_batch b = {}; int64_t nmax;
    return ew_build(tasks, n, b, nmax);
}
extern "C" int b200_op_rope_cpy_batch(const b200_ew_desc * tasks, int n, void * stream) {
    ew_batch b = {}; int64_t nmax;
    const int rc = ew_build(tasks, n, b, nmax);
    if (rc) return rc;
    if (nmax == 0) return 0;
    // a prefill batch: enough (token, pair) positions that the per-head-group mapping fills the GPU
    int64_t wide = 0; bool all_wide = true;
    for (int i = 0; i < n; ++i) {
        const ew_task & t = b.t[i];
        if (t.kind != 2 && (t.src.ne[1] < 4 || t.src.ne[0] / 2 * t.src.ne[2] * t.src.ne[3] < 8192)) all_wide = false;
        wide = std::max(wide, t.kind == 2 && !t.vec4 ? t.n : t.nwide);
    }
    if (all_wide && nmax >= 16384) return launch(ew_batch_wide_kernel, dim3(grid_for(wide, 256), (unsigned)n), 256, 0, (cudaStream_t)stream, b);
    return launch(ew_batch_kernel, dim3(grid_for(nmax, 256), (unsigned)n), 256, 0, (cudaStream_t)stream, b);
}
