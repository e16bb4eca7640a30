// fused.cu -- fusions of adjacent graph nodes.  A decode step of an 8B model is ~800 kernels of a few microseconds each and
// a prefill pass spends a quarter of its time in element-wise glue; these kernels do the work of 3-4 nodes in one launch and
// hand the next mat-mul its activations in the form it reads (q8_1 blocks for the GEMV, bf16 rows for the MMQ).  Every
// intermediate tensor of the original nodes is still written, with the same arithmetic in the same order as the separate
// kernels (glue.cu), so results are bit-identical with and without fusion and nothing needs to be known about other
// consumers of the intermediates.
//   b200_op_add_rms_norm_mul : [ADD] -> RMS_NORM -> MUL(weight)   (+ activations of the result for the next mat-mul)
//   b200_op_unary_mul        : UNARY(silu|gelu) -> MUL(other)     (+ activations of the result)
// Reference: the same chains are fused in ggml-cuda (RMS_NORM+MUL in norm.cu / ggml-cuda.cu graph fusion; GLU kernels in
// unary.cu); semantics of each step follow ggml-cpu/ops.cpp (rms_norm 3285-3312) and unary-ops.cpp (silu, gelu).
#include "mmvq_r.cuh"        // sumsq4 / rms_scale: the norm arithmetic shared with the GEMV's fused prologue
#include "../../include/b200_ops.h"

#include <algorithm>

namespace b200 {

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi);
// one CTA per row, thread t holds float4 t, t + blockDim, ... of the row (8 consecutive threads = one 32-block)
template <int VPT>
__global__ void __launch_bounds__(1024) add_rms_norm_mul_kernel(const float * __restrict__ a, int64_t as, const float * __restrict__ b, int64_t bs,
                                                                float * __restrict__ sum, int64_t ss, float * __restrict__ norm, int64_t ns,
                                                                const float * __restrict__ w, float * __restrict__ out, int64_t os,
                                                                uint8_t * __restrict__ xq, int xkind, int K, float eps) {
    pdl_trigger(); pdl_wait();
    const int r = blockIdx.x, n4 = K >> 2;
    const float4 * a4 = reinterpret_cast<const float4 *>(a + (int64_t)r * as);
    float4 v[VPT];
#pragma unroll
    for (int k = 0; k < VPT; ++k) { const int i = threadIdx.x + k * blockDim.x; v[k] = i < n4 ? a4[i] : make_float4(0.f, 0.f, 0.f, 0.f); }
    if (b) {
        const float4 * b4 = reinterpret_cast<const float4 *>(b + (int64_t)r * bs);
        float4 * s4 = reinterpret_cast<float4 *>(sum + (int64_t)r * ss);
#pragma unroll
        for (int k = 0; k < VPT; ++k) {
            const int i = threadIdx.x + k * blockDim.x;
            if (i < n4) { const float4 y = b4[i]; v[k] = make_float4(v[k].x + y.x, v[k].y + y.y, v[k].z + y.z, v[k].w + y.w); s4[i] = v[k]; }
        }
    }
    float s = 0.0f;
#pragma unroll
    for (int k = 0; k < VPT; ++k) s += sumsq4(v[k]);
    s = warp_sum(s);
    __shared__ float part[32];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
    __syncthreads();
    float tot = 0.0f;
    for (int wi = 0; wi < (int)(blockDim.x >> 5); ++wi) tot += part[wi];
    const float scale = rms_scale(tot, K, eps);
    const float4 * w4 = reinterpret_cast<const float4 *>(w);
    float4 * n4p = reinterpret_cast<float4 *>(norm + (int64_t)r * ns);
    float4 * o4 = reinterpret_cast<float4 *>(out + (int64_t)r * os);
    const uint32_t acs = (uint32_t)((K + 255) / 256) * ACT_ITEM;
#pragma unroll
    for (int k = 0; k < VPT; ++k) {
        const int i = threadIdx.x + k * blockDim.x;
        float4 y = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i < n4) {
            const float4 nv = make_float4(v[k].x * scale, v[k].y * scale, v[k].z * scale, v[k].w * scale);
            const float4 ww = w4[i];
            y = make_float4(nv.x * ww.x, nv.y * ww.y, nv.z * ww.z, nv.w * ww.w);
            n4p[i] = nv; o4[i] = y;
        }
        if (xq && xkind == 0) quant_block_store(y, i >> 3, threadIdx.x & 7, i < n4, xq + (size_t)r * acs);
        if (xq && xkind == 1 && i < n4) reinterpret_cast<uint2 *>(xq + (size_t)r * K * 2)[i] = make_uint2(pack_bf16x2(y.x, y.y), pack_bf16x2(y.z, y.w));
    }
}

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
    uint32_t d; asm("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(d) : "f"(hi), "f"(lo)); return d;
}
template <int OP> __device__ __forceinline__ float act_fn(float x) {
    if constexpr (OP == B200_UNARY_SILU) return silu_fn(x);
    else { const float c = 0.044715f, s = 0.79788456080286535587989211986876f; return 0.5f * x * (1.0f + tanhf(s * x * (1.0f + c * x * x))); }
}
// rows of K floats, flat over (row, float4); 8 consecutive threads = one 32-block of one row (K % 32 == 0)
template <int OP>
__global__ void __launch_bounds__(256) unary_mul_kernel(const float * __restrict__ g, int64_t gs, float * __restrict__ act, int64_t as,
                                                        const float * __restrict__ u, int64_t us, float * __restrict__ out, int64_t os,
                                                        uint8_t * __restrict__ xq, int xkind, int K, int M) {
    pdl_trigger(); pdl_wait();
    const int n4 = K >> 2;
    const int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool in = idx < (int64_t)M * n4;
    const int r = in ? (int)(idx / n4) : 0, i = in ? (int)(idx - (int64_t)r * n4) : 0;
    float4 y = make_float4(0.f, 0.f, 0.f, 0.f);
    if (in) {
        const float4 x = reinterpret_cast<const float4 *>(g + (int64_t)r * gs)[i];
        const float4 a = make_float4(act_fn<OP>(x.x), act_fn<OP>(x.y), act_fn<OP>(x.z), act_fn<OP>(x.w));
        const float4 uu = reinterpret_cast<const float4 *>(u + (int64_t)r * us)[i];
        y = make_float4(a.x * uu.x, a.y * uu.y, a.z * uu.z, a.w * uu.w);
        reinterpret_cast<float4 *>(act + (int64_t)r * as)[i] = a;
        reinterpret_cast<float4 *>(out + (int64_t)r * os)[i] = y;
    }
    if (xq && xkind == 0) {
        const uint32_t acs = (uint32_t)((K + 255) / 256) * ACT_ITEM;
        quant_block_store(y, i >> 3, threadIdx.x & 7, in, xq + (size_t)r * acs);
    }
    if (xq && xkind == 1 && in) reinterpret_cast<uint2 *>(xq + (size_t)r * K * 2)[i] = make_uint2(pack_bf16x2(y.x, y.y), pack_bf16x2(y.z, y.w));
}

template <typename... KArgs, typename... Args>
static int flaunch(void (*kern)(KArgs...), dim3 grid, dim3 block, cudaStream_t st, Args... args) {
    cudaLaunchConfig_t lc = {};
    lc.gridDim = grid; lc.blockDim = block; lc.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = pdl_default() ? 1 : 0;
    B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, kern, KArgs(args)...));
    ++g_launches;
    return 0;
}

// rows of t as a [M][K] f32 matrix with 16-byte aligned rows: M = product of the outer dims when they are packed
static bool rows_of(const b200_tensor * t, int K, int M, int64_t * stride) {
    if (t->type != T_F32 || t->ne[0] != K || t->nb[0] != 4 || t->ne[1] * t->ne[2] * t->ne[3] != M) return false;
    if (((uintptr_t)t->data | t->nb[1]) & 15) return false;
    if (t->ne[2] != 1 && t->nb[2] != t->nb[1] * t->ne[1]) return false;
    if (t->ne[3] != 1 && t->nb[3] != t->nb[2] * t->ne[2]) return false;
    *stride = t->nb[1] / 4;
    return true;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_op_add_rms_norm_mul(const b200_tensor * a, const b200_tensor * b, const b200_tensor * sum, const b200_tensor * norm,
                                        const b200_tensor * w, const b200_tensor * out, float eps, void * xq, int xq_kind, void * stream) {
    const int K = (int)a->ne[0];
    const int64_t M64 = a->ne[1] * a->ne[2] * a->ne[3];
    if (K <= 0 || K % 32 || K > 16384 || M64 < 1 || M64 > 65535 || (xq && xq_kind == 0 && M64 > 8) || (xq_kind != 0 && xq_kind != 1)) return -1;
    const int M = (int)M64;
    if (xq && ((uintptr_t)xq & 15)) return -1;
    int64_t as, bs = 0, ss = 0, ns, os;
    if (!rows_of(a, K, M, &as) || !rows_of(norm, K, M, &ns) || !rows_of(out, K, M, &os)) return -1;
    if (b && (!sum || !rows_of(b, K, M, &bs) || !rows_of(sum, K, M, &ss))) return -1;
    if (w->type != T_F32 || w->ne[0] != K || w->ne[1] * w->ne[2] * w->ne[3] != 1 || w->nb[0] != 4 || ((uintptr_t)w->data & 15)) return -1;
    const int n4 = K / 4;
    // few rows (decode): as many threads as float4s, one each (pure latency); many rows (prefill): 256 threads, up to 4 each
    const int threads = M <= 8 || n4 > 1024 ? std::min(1024, (n4 + 31) / 32 * 32) : std::min(256, (n4 + 31) / 32 * 32);
    cudaStream_t st = (cudaStream_t)stream;
    const float * bp = b ? (const float *)b->data : nullptr;
    float * sp = b ? (float *)sum->data : nullptr;
#define B200_NORM_ARGS (const float *)a->data, as, bp, bs, sp, ss, (float *)norm->data, ns, (const float *)w->data, (float *)out->data, os, (uint8_t *)xq, xq_kind, K, eps
    if (n4 <= threads) return flaunch(add_rms_norm_mul_kernel<1>, M, threads, st, B200_NORM_ARGS);
    if (n4 <= 2 * threads) return flaunch(add_rms_norm_mul_kernel<2>, M, threads, st, B200_NORM_ARGS);
    return flaunch(add_rms_norm_mul_kernel<4>, M, threads, st, B200_NORM_ARGS);
#undef B200_NORM_ARGS
}

extern "C" int b200_op_unary_mul(int op, const b200_tensor * g, const b200_tensor * act, const b200_tensor * u, const b200_tensor * out,
                                 void * xq, int xq_kind, void * stream) {
    const int K = (int)g->ne[0];
    const int64_t M64 = g->ne[1] * g->ne[2] * g->ne[3];
    if (K <= 0 || K % 32 || M64 < 1 || M64 > 65535 || (op != B200_UNARY_SILU && op != B200_UNARY_GELU)) return -1;
    if ((xq && xq_kind == 0 && M64 > 8) || (xq_kind != 0 && xq_kind != 1)) return -1;
    const int M = (int)M64;
    int64_t gs, as, us, os;
    if (!rows_of(g, K, M, &gs) || !rows_of(act, K, M, &as) || !rows_of(u, K, M, &us) || !rows_of(out, K, M, &os)) return -1;
    const int64_t n = (int64_t)M * (K / 4);
    const unsigned grid = (unsigned)((n + 255) / 256);
    cudaStream_t st = (cudaStream_t)stream;
    if (op == B200_UNARY_SILU)
        return flaunch(unary_mul_kernel<B200_UNARY_SILU>, grid, 256, st, (const float *)g->data, gs, (float *)act->data, as, (const float *)u->data, us,
                       (float *)out->data, os, (uint8_t *)xq, xq_kind, K, M);
    return flaunch(unary_mul_kernel<B200_UNARY_GELU>, grid, 256, st, (const float *)g->data, gs, (float *)act->data, as, (const float *)u->data, us,
                   (float *)out->data, os, (uint8_t *)xq, xq_kind, K, M);
}
