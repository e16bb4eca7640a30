// matmul.cu -- GGML_OP_MUL_MAT / MUL_MAT_ID dispatch (reference ggml-cuda/ggml-cuda.cu:1936-2005,
// 2007-2148; mmvq.cu:491-561) onto the B200 kernels, plus a plain tiled kernel for f16/bf16/f32
// weights (the K.Q and V.P products of the non-flash attention path).
#include "mmvq_units.cuh"
#include "../../include/b200_ops.h"

#include <algorithm>
#include <cstdlib>
#include <cuda_bf16.h>

namespace b200 {

// ---- dense src0 (f32 / f16 / bf16) x src1 (f32 / f16): 32x32 output tile per CTA, K in steps of 32 ----
template <typename T> __device__ __forceinline__ float ld_f(const char * p);
template <> __device__ __forceinline__ float ld_f<float>(const char * p) { return *(const float *)p; }
template <> __device__ __forceinline__ float ld_f<__half>(const char * p) { return __half2float(*(const __half *)p); }
template <> __device__ __forceinline__ float ld_f<__nv_bfloat16>(const char * p) { return __bfloat162float(*(const __nv_bfloat16 *)p); }

struct mm_params {
    const char * a; const char * b; char * d;
    int64_t nba0, nba1, nba2, nba3, nbb0, nbb1, nbb2, nbb3, nbd1, nbd2, nbd3;
    int K, N, M, ne12, r2, r3;
};
template <typename TA, typename TB>
__global__ void __launch_bounds__(256) mm_dense_kernel(const mm_params p) {
    pdl_trigger(); pdl_wait();
    __shared__ float sa[32][33], sb[32][33];
    const int i12 = blockIdx.z % p.ne12, i13 = blockIdx.z / p.ne12;
    const char * A = p.a + (int64_t)(i12 / p.r2) * p.nba2 + (int64_t)(i13 / p.r3) * p.nba3;
    const char * B = p.b + (int64_t)i12 * p.nbb2 + (int64_t)i13 * p.nbb3;
    const int n0 = blockIdx.x * 32, m0 = blockIdx.y * 32;
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;          // 32 x 8
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    for (int k0 = 0; k0 < p.K; k0 += 32) {
        for (int r = ty; r < 32; r += 8) {
            const int k = k0 + tx;
            sa[r][tx] = (n0 + r < p.N && k < p.K) ? ld_f<TA>(A + (int64_t)(n0 + r) * p.nba1 + (int64_t)k * p.nba0) : 0.0f;
            sb[r][tx] = (m0 + r < p.M && k < p.K) ? ld_f<TB>(B + (int64_t)(m0 + r) * p.nbb1 + (int64_t)k * p.nbb0) : 0.0f;
        }
        __syncthreads();
#pragma unroll 8
        for (int k = 0; k < 32; ++k) {
            const float av = sa[tx][k];
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[j] = fmaf(av, sb[ty + 8 * j][k], acc[j]);
        }
        __syncthreads();
    }
    for (int j = 0; j < 4; ++j) {
        const int n = n0 + tx, m = m0 + ty + 8 * j;
        if (n < p.N && m < p.M) *(float *)(p.d + (int64_t)m * p.nbd1 + (int64_t)i12 * p.nbd2 + (int64_t)i13 * p.nbd3 + (int64_t)n * 4) = acc[j];
    }
}

// ---- few token columns (M <= 8): one CTA per weight row, K spread over the threads ----
// (the router of a MoE layer: [n_embd] x [n_expert] in f32; K.Q and V.P of the non-flash attention path at decode)
// NT threads per row: the MoE router (K = 4096, 8 rows, 8 CTAs in all) is pure latency -- with 128 threads each lane walked 32
// dependent-looking iterations (21 us per call, 11 % of a Mixtral decode step); 512 threads and four loads in flight: ~4 us.
template <typename TA, typename TB, int NT>
__global__ void __launch_bounds__(NT) mv_dense_kernel(const mm_params p) {
    pdl_trigger(); pdl_wait();
    const int i12 = blockIdx.z % p.ne12, i13 = blockIdx.z / p.ne12;
    const char * A = p.a + (int64_t)(i12 / p.r2) * p.nba2 + (int64_t)(i13 / p.r3) * p.nba3 + (int64_t)blockIdx.x * p.nba1;
    const int m0 = 8 * blockIdx.y, mc = min(8, p.M - m0);        // this CTA's chunk of (up to 8) token columns
    const char * B = p.b + (int64_t)i12 * p.nbb2 + (int64_t)i13 * p.nbb3 + (int64_t)m0 * p.nbb1;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll 4
    for (int k = threadIdx.x; k < p.K; k += NT) {
        const float av = ld_f<TA>(A + (int64_t)k * p.nba0);
#pragma unroll
        for (int m = 0; m < 8; ++m) if (m < mc) acc[m] = fmaf(av, ld_f<TB>(B + (int64_t)m * p.nbb1 + (int64_t)k * p.nbb0), acc[m]);
    }
    __shared__ float part[NT / 32][8];
#pragma unroll
    for (int m = 0; m < 8; ++m) {
        float v = acc[m];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5][m] = v;
    }
    __syncthreads();
    if (threadIdx.x < mc) {
        float v = 0.0f;
#pragma unroll
        for (int w = 0; w < NT / 32; ++w) v += part[w][threadIdx.x];
        *(float *)(p.d + (int64_t)(m0 + threadIdx.x) * p.nbd1 + (int64_t)i12 * p.nbd2 + (int64_t)i13 * p.nbd3 + (int64_t)blockIdx.x * 4) = v;
    }
}

template <typename TA, typename TB>
static int mm_dense_launch(const mm_params & p, int nz, cudaStream_t st) {
    // few token columns, or a few rows x a modest batch (the MoE router of a batched decode step: 8 rows x 9 .. 64 tokens went to the
    // 32 x 32 tile kernel as ONE CTA, 276 us per call -- most of the gap between 8 and 12 sequences in llama-batched-bench)
    if (p.M <= 8 || (p.N <= 64 && p.M <= 512)) {
        cudaLaunchConfig_t lc = {};
        const bool wide = p.K >= 2048 && (int64_t)p.N * nz * ((p.M + 7) / 8) <= 1024;       // few long rows: more threads per row
        lc.gridDim = dim3(p.N, (p.M + 7) / 8, nz); lc.blockDim = dim3(wide ? 512 : 128); lc.stream = st;
        cudaLaunchAttribute at[1];
        at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
        at[0].val.programmaticStreamSerializationAllowed = 1;
        lc.attrs = at; lc.numAttrs = pdl_default() ? 1 : 0;
        if (wide) B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, mv_dense_kernel<TA, TB, 512>, p));
        else B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, mv_dense_kernel<TA, TB, 128>, p));
        ++g_launches;
        return 0;
    }
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3((p.N + 31) / 32, (p.M + 31) / 32, nz); lc.blockDim = dim3(256); lc.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = pdl_default() ? 1 : 0;
    B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, mm_dense_kernel<TA, TB>, p));
    ++g_launches;
    return 0;
}

static bool is_quant(int t) { return b200_mmvq_supported_type(t) != 0; }
static bool is_dense(int t) { return t == T_F32 || t == T_F16 || t == T_BF16; }

static int mmq_min_m() { static const int v = [] { const char * e = getenv("B200_MMQ_MIN_M"); return e ? atoi(e) : 9; }(); return v; }
// the tensor-core path takes quantised weights x one plane of more than 8 token columns
static bool use_mmq(const b200_tensor * a, const b200_tensor * b) {
    return is_quant(a->type) && b->ne[2] * b->ne[3] == 1 && b->ne[1] >= mmq_min_m() && b->ne[1] > 8 &&
           b200_mul_mat_q_supported(a->type, a->data, a->nb[1], (int)a->ne[0]);
}
static int xq_min_k() { static const int v = [] { const char * e = getenv("B200_XQ_MIN_K"); return e ? atoi(e) : 8192; }(); return v; }

}  // namespace b200

using namespace b200;

extern "C" int b200_mul_mat_supported(const b200_tensor * a, const b200_tensor * b) {
    if (a->ne[0] != b->ne[0]) return 0;
    if (a->ne[2] <= 0 || a->ne[3] <= 0 || b->ne[2] % a->ne[2] || b->ne[3] % a->ne[3]) return 0;
    if (is_quant(a->type)) {
        if (b->type != T_F32 || b->nb[0] != 4) return 0;
        if (a->ne[0] % block_elems(a->type) || a->nb[0] != block_bytes(a->type)) return 0;
        if ((b->nb[1] | b->nb[2] | b->nb[3]) & 3) return 0;
        if (a->ne[0] > 61440) return 0;                 // quantised activations of one column must fit shared memory
        return 1;
    }
    if (is_dense(a->type)) return b->type == T_F32 || b->type == T_F16;
    return 0;
}

extern "C" int64_t b200_mul_mat_scratch_bytes(const b200_tensor * a, const b200_tensor * b) {
    if (use_mmq(a, b)) return b200_mul_mat_q_scratch_bytes((int)a->ne[0], (int)b->ne[1]);
    if (!is_quant(a->type) || b->ne[2] * b->ne[3] != 1 || a->ne[0] < xq_min_k()) return 0;
    return b200_quantized_acts_bytes((int)a->ne[0], (int)b->ne[1]);
}

extern "C" int b200_op_mul_mat(const b200_tensor * a, const b200_tensor * b, const b200_tensor * dst, void * scratch, int64_t scratch_bytes,
                               void * stream) {
    if (!b200_mul_mat_supported(a, b) || dst->type != T_F32 || dst->nb[0] != 4) return -1;
    const int K = (int)a->ne[0], N = (int)a->ne[1], M = (int)b->ne[1];
    if (dst->ne[0] != N || dst->ne[1] != M || dst->ne[2] != b->ne[2] || dst->ne[3] != b->ne[3]) return -1;
    if (K == 0 || N == 0 || M == 0 || b->ne[2] * b->ne[3] == 0) return 0;
    if (use_mmq(a, b) && scratch && scratch_bytes >= b200_mul_mat_q_scratch_bytes(K, M))
        return b200_mul_mat_q(a->type, a->data, a->nb[1], (const float *)b->data, b->nb[1] / 4, (float *)dst->data, dst->nb[1] / 4,
                              N, K, M, scratch, scratch_bytes, stream);
    if (is_quant(a->type)) {
        b200_mmvq_args g = {};
        g.type = a->type; g.W = a->data; g.nb1 = a->nb[1];
        g.X = (const float *)b->data; g.xs = b->nb[1] / 4;
        g.dst = (float *)dst->data; g.ds = dst->nb[1] / 4;
        g.N = N; g.K = K; g.M = M;
        g.ny = (int)b->ne[2]; g.nz = (int)b->ne[3];
        g.w_sy = a->nb[2]; g.w_sz = a->nb[3]; g.w_div_y = (int)(b->ne[2] / a->ne[2]); g.w_div_z = (int)(b->ne[3] / a->ne[3]);
        g.x_sy = b->nb[2] / 4; g.x_sz = b->nb[3] / 4; g.d_sy = dst->nb[2] / 4; g.d_sz = dst->nb[3] / 4;
        g.pdl = -1;
        const int64_t need = use_mmq(a, b) ? 0 : b200_mul_mat_scratch_bytes(a, b);
        if (need > 0 && scratch && scratch_bytes >= need) {
            const int rc = b200_quantize_acts(g.X, g.xs, scratch, K, M, stream);
            if (rc) return rc;
            g.Xq = scratch;
        }
        return b200_mul_mat_vec_q_ex(&g, stream);
    }
    mm_params p = {};
    p.a = (const char *)a->data; p.b = (const char *)b->data; p.d = (char *)dst->data;
    p.nba0 = a->nb[0]; p.nba1 = a->nb[1]; p.nba2 = a->nb[2]; p.nba3 = a->nb[3];
    p.nbb0 = b->nb[0]; p.nbb1 = b->nb[1]; p.nbb2 = b->nb[2]; p.nbb3 = b->nb[3];
    p.nbd1 = dst->nb[1]; p.nbd2 = dst->nb[2]; p.nbd3 = dst->nb[3];
    p.K = K; p.N = N; p.M = M; p.ne12 = (int)b->ne[2]; p.r2 = (int)(b->ne[2] / a->ne[2]); p.r3 = (int)(b->ne[3] / a->ne[3]);
    const int nz = (int)(b->ne[2] * b->ne[3]);
    if (nz > 65535) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    const bool bf = b->type == T_F32;
    switch (a->type) {
        case T_F32:  return bf ? mm_dense_launch<float, float>(p, nz, st) : mm_dense_launch<float, __half>(p, nz, st);
        case T_F16:  return bf ? mm_dense_launch<__half, float>(p, nz, st) : mm_dense_launch<__half, __half>(p, nz, st);
        case T_BF16: return bf ? mm_dense_launch<__nv_bfloat16, float>(p, nz, st) : mm_dense_launch<__nv_bfloat16, __half>(p, nz, st);
    }
    return -1;
}

// batches: rows sorted by expert + tensor-core MMQ (mmq.cu); a few tokens: one GEMV per (token, expert) pair
static int mmq_id_min_t() { static const int v = [] { const char * e = getenv("B200_MMQ_ID_MIN_T"); return e ? atoi(e) : 3; }(); return v; }
static bool use_mmq_id(const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids) {
    if (ids->ne[1] < mmq_id_min_t() || as->nb[2] != as->ne[1] * as->nb[1] || as->ne[2] > 4096 || (ids->nb[1] & 3) || ids->nb[0] != 4) return false;
    return b200_mul_mat_q_supported(as->type, as->data, as->nb[1], (int)as->ne[0]) != 0 && (b->nb[1] & 3) == 0 && (b->nb[2] & 3) == 0;
}
// one token, K-quant experts: the register-resident GEMV takes the (up to 3) chosen experts as a group launch; it needs the
// activation columns quantised first (scratch: one 400-byte item per 256 values and column)
static bool use_mmvqr_id(const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids) {
    static const bool on = [] { const char * e = getenv("B200_MMVQR_ID"); return !e || atoi(e) != 0; }();
    return on && is_quant(as->type) && ids->ne[1] == 1 && ids->ne[0] <= 3 && as->ne[0] % 256 == 0 && (b->nb[1] & 15) == 0 &&
           ((uintptr_t)b->data & 15) == 0 && (as->nb[2] & 15) == 0;
}
extern "C" int64_t b200_mul_mat_id_scratch_bytes(const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids) {
    if (!is_quant(as->type)) return 0;
    if (use_mmvqr_id(as, b, ids)) return b->ne[1] * (as->ne[0] / 256) * ACT_ITEM + 256;
    if (!use_mmq_id(as, b, ids)) return 0;
    return b200_mul_mat_id_q_scratch_bytes((int)as->ne[0], (int)ids->ne[0], (int)ids->ne[1], (int)as->ne[2]);
}

extern "C" int b200_op_mul_mat_id(const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids, const b200_tensor * dst,
                                  void * scratch, int64_t scratch_bytes, void * stream) {
    if (!is_quant(as->type) || b->type != T_F32 || ids->type != T_I32 || dst->type != T_F32) return -1;
    if (as->ne[0] != b->ne[0] || b->nb[0] != 4 || dst->nb[0] != 4 || as->nb[0] != block_bytes(as->type)) return -1;
    if (as->ne[0] % block_elems(as->type) || as->ne[3] != 1 || b->ne[3] != 1) return -1;
    const int K = (int)as->ne[0], N = (int)as->ne[1];
    const int n_used = (int)ids->ne[0], T = (int)ids->ne[1];
    if (dst->ne[0] != N || dst->ne[1] != n_used || dst->ne[2] != T || b->ne[2] != T || b->ne[1] <= 0 || n_used % b->ne[1]) {
        if (!(b->ne[1] == 1 || b->ne[1] == n_used)) return -1;
    }
    if (T > 65535 || n_used > 65535) return -1;
    if (K == 0 || N == 0 || n_used == 0 || T == 0) return 0;
    if (use_mmq_id(as, b, ids) && scratch && scratch_bytes >= b200_mul_mat_id_q_scratch_bytes(K, n_used, T, (int)as->ne[2])) {
        const int rc = b200_mul_mat_id_q(as->type, as->data, as->nb[1], (int)as->ne[2], (const float *)b->data, b->nb[1] / 4, b->nb[2] / 4, (int)b->ne[1],
                                         (const int32_t *)ids->data, ids->nb[1] / 4, n_used, T, (float *)dst->data, dst->nb[1] / 4, dst->nb[2] / 4,
                                         N, K, scratch, scratch_bytes, stream);
        if (rc != -1) return rc;
    }
    b200_mmvq_args g = {};
    g.type = as->type; g.W = as->data; g.nb1 = as->nb[1];
    g.X = (const float *)b->data; g.xs = K;
    g.dst = (float *)dst->data; g.ds = N;
    g.N = N; g.K = K; g.M = 1;
    g.ny = n_used; g.nz = T;
    g.w_sy = as->nb[2];
    g.x_sy = b->nb[1] / 4; g.x_sz = b->nb[2] / 4; g.x_mod_y = (int)b->ne[1];
    g.d_sy = dst->nb[1] / 4; g.d_sz = dst->nb[2] / 4;
    g.ids = (const int32_t *)ids->data; g.ids_sz = ids->nb[1] / 4;
    g.pdl = -1;
    if (use_mmvqr_id(as, b, ids) && scratch && scratch_bytes >= b->ne[1] * (K / 256) * ACT_ITEM + 256) {
        // quantise the activation column(s) once, then one group launch of the register-resident kernel over the chosen experts
        void * xq = (void *)(((uintptr_t)scratch + 255) & ~(uintptr_t)255);
        b200_mmvq_args h = g;
        h.Xq = xq;
        if (b200_mmvq_fusable(&h, 0, nullptr, nullptr)) {
            const int rc = b200_quantize_acts((const float *)b->data, b->nb[1] / 4, xq, K, (int)b->ne[1], stream);
            if (rc) return rc;
            return b200_mul_mat_vec_q_ex(&h, stream);
        }
    }
    return b200_mul_mat_vec_q_ex(&g, stream);
}
