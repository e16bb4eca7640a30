// glue_rows.cu -- row-wise operators: GET_ROWS, SOFT_MAX, ARGSORT, SUM_ROWS.
// Semantics follow the reference CPU operators (ggml/src/ggml-cpu/ops.cpp), cited per kernel.
#include "dequant.cuh"
#include "../../include/b200_ops.h"

#include <algorithm>
#include <cmath>
#include <cuda_bf16.h>

namespace b200 {

struct rtens { char * d; int64_t ne[4]; int64_t nb[4]; };
static rtens rmk(const b200_tensor * t) {
    rtens r; r.d = (char *)t->data;
    for (int i = 0; i < 4; ++i) { r.ne[i] = t->ne[i]; r.nb[i] = t->nb[i]; }
    return r;
}
template <typename... KArgs, typename... Args>
static int rlaunch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
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

// ---- get_rows (reference ggml-cpu/ops.cpp:4375-4399): dst[:, i10, i11, i12] = src[:, ids[i10,i11,i12], i11, i12] ----
template <int TYPE> __device__ __forceinline__ float row_elem(const char * row, int64_t e) {
    if constexpr (TYPE == T_F32) return ((const float *)row)[e];
    else if constexpr (TYPE == T_F16) return __half2float(((const __half *)row)[e]);
    else if constexpr (TYPE == T_BF16) return __bfloat162float(((const __nv_bfloat16 *)row)[e]);
    else return dequant_elem<TYPE>((const uint8_t *)row, e);
}
template <int TYPE>
__global__ void __launch_bounds__(256) get_rows_kernel(const rtens src, const rtens ids, const rtens dst) {
    pdl_trigger(); pdl_wait();
    const int64_t r = blockIdx.x;                                    // destination row
    const int64_t i10 = r % ids.ne[0], i11 = (r / ids.ne[0]) % ids.ne[1], i12 = r / (ids.ne[0] * ids.ne[1]);
    const int64_t i01 = *(const int32_t *)(ids.d + i10 * ids.nb[0] + i11 * ids.nb[1] + i12 * ids.nb[2]);
    const char * row = src.d + i01 * src.nb[1] + i11 * src.nb[2] + i12 * src.nb[3];
    float * out = (float *)(dst.d + i10 * dst.nb[1] + i11 * dst.nb[2] + i12 * dst.nb[3]);
    for (int64_t e = threadIdx.x; e < src.ne[0]; e += blockDim.x) out[e] = row_elem<TYPE>(row, e);
}

// ---- soft_max (reference ggml-cpu/ops.cpp:4731-4826): one CTA per row, row cached in shared memory ----
template <typename MT>
__global__ void __launch_bounds__(1024) soft_max_kernel(const float * __restrict__ x, const MT * __restrict__ mask, float * __restrict__ y,
                                                        int nc, int64_t ne01, int64_t ne02, float scale, float max_bias, float m0, float m1,
                                                        unsigned n_head_log2) {
    pdl_trigger(); pdl_wait();
    extern __shared__ float row[];
    __shared__ float red[32];
    const int64_t i1 = blockIdx.x;
    const unsigned h = (unsigned)((i1 / ne01) % ne02);
    const float slope = max_bias > 0.0f ? (h < n_head_log2 ? powf(m0, (float)(h + 1)) : powf(m1, (float)(2 * (h - n_head_log2) + 1))) : 1.0f;
    const float * sp = x + i1 * nc;
    const MT * mp = mask ? mask + (i1 % ne01) * nc : nullptr;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, nw = blockDim.x >> 5;
    float mx = -INFINITY;
    for (int i = tid; i < nc; i += blockDim.x) {
        float v = sp[i] * scale;
        if (mp) v += slope * (float)mp[i];
        row[i] = v;
        mx = fmaxf(mx, v);
    }
    mx = warp_max(mx);
    if (lane == 0) red[warp] = mx;
    __syncthreads();
    mx = red[0];
    for (int w = 1; w < nw; ++w) mx = fmaxf(mx, red[w]);
    __syncthreads();
    float sum = 0.0f;
    for (int i = tid; i < nc; i += blockDim.x) { const float e = expf(row[i] - mx); row[i] = e; sum += e; }
    sum = warp_sum(sum);
    if (lane == 0) red[warp] = sum;
    __syncthreads();
    float tot = 0.0f;
    for (int w = 0; w < nw; ++w) tot += red[w];
    const float inv = 1.0f / tot;
    for (int i = tid; i < nc; i += blockDim.x) y[i1 * nc + i] = row[i] * inv;
}

// ---- argsort (reference ggml-cpu/ops.cpp ggml_compute_forward_argsort_f32): bitonic sort of one row per CTA ----
__global__ void __launch_bounds__(1024) argsort_kernel(const float * __restrict__ x, int32_t * __restrict__ dst, int ncols, int npad, int desc) {
    pdl_trigger(); pdl_wait();
    extern __shared__ int idx[];
    const int row = blockIdx.x, col = threadIdx.x;
    const float * xr = x + (int64_t)row * ncols;
    if (col < npad) idx[col] = col;
    __syncthreads();
    for (int k = 2; k <= npad; k *= 2) {
        for (int j = k / 2; j > 0; j /= 2) {
            const int ixj = col ^ j;
            if (col < npad && ixj > col) {
                const int a = idx[col], b = idx[ixj];
                // padding indices (>= ncols) sort to the end in either order
                bool swap;
                const bool up = (col & k) == 0;
                if (a >= ncols) swap = up ? (b < ncols) : false;
                else if (b >= ncols) swap = up ? false : true;
                else {
                    const bool gt = desc ? xr[a] < xr[b] : xr[a] > xr[b];
                    const bool lt = desc ? xr[a] > xr[b] : xr[a] < xr[b];
                    swap = up ? gt : lt;
                }
                if (swap) { idx[col] = b; idx[ixj] = a; }
            }
            __syncthreads();
        }
    }
    if (col < ncols) dst[(int64_t)row * ncols + col] = idx[col];
}

// ---- sum_rows: one warp per row ----
__global__ void __launch_bounds__(256) sum_rows_kernel(const rtens src, const rtens dst, int64_t nrows) {
    pdl_trigger(); pdl_wait();
    const int64_t r = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (r >= nrows) return;
    const int64_t i1 = r % src.ne[1], i2 = (r / src.ne[1]) % src.ne[2], i3 = r / (src.ne[1] * src.ne[2]);
    const float * x = (const float *)(src.d + i1 * src.nb[1] + i2 * src.nb[2] + i3 * src.nb[3]);
    float s = 0.0f;
    for (int64_t i = threadIdx.x & 31; i < src.ne[0]; i += 32) s += x[i];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) *(float *)(dst.d + i1 * dst.nb[1] + i2 * dst.nb[2] + i3 * dst.nb[3]) = s;
}

static bool contig_f(const b200_tensor * t, int esz) {
    int64_t s = esz;
    for (int i = 0; i < 4; ++i) { if (t->ne[i] != 1 && t->nb[i] != s) return false; s *= t->ne[i]; }
    return true;
}

}  // namespace b200

using namespace b200;

#define B200_GR_CASE(T) case T: return rlaunch(get_rows_kernel<T>, (unsigned)nr, threads, 0, st, rmk(src), rmk(ids), rmk(dst));
extern "C" int b200_op_get_rows(const b200_tensor * src, const b200_tensor * ids, const b200_tensor * dst, void * stream) {
    if (ids->type != T_I32 || dst->type != T_F32 || dst->nb[0] != 4 || src->nb[0] != block_bytes(src->type)) return -1;
    if (dst->ne[0] != src->ne[0]) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    const int64_t nr = ids->ne[0] * ids->ne[1] * ids->ne[2];
    if (nr == 0 || src->ne[0] == 0) return 0;
    const int threads = src->ne[0] >= 1024 ? 256 : 64;
    switch (src->type) {
        B200_GR_CASE(T_F32) B200_GR_CASE(T_F16) B200_GR_CASE(T_BF16) B200_GR_CASE(T_Q4_0) B200_GR_CASE(T_Q8_0)
        B200_GR_CASE(T_Q4_K) B200_GR_CASE(T_Q5_K) B200_GR_CASE(T_Q6_K) B200_GR_CASE(T_Q4_1) B200_GR_CASE(T_Q5_0) B200_GR_CASE(T_Q5_1)
    }
    return -1;
}

extern "C" int b200_op_soft_max(const b200_tensor * src, const b200_tensor * mask, const b200_tensor * dst, float scale, float max_bias,
                                void * stream) {
    if (src->type != T_F32 || dst->type != T_F32 || !contig_f(src, 4) || !contig_f(dst, 4)) return -1;
    if (mask && ((mask->type != T_F16 && mask->type != T_F32) || !contig_f(mask, block_bytes(mask->type)) || mask->ne[0] != src->ne[0])) return -1;
    const int nc = (int)src->ne[0];
    const int64_t nr = src->ne[1] * src->ne[2] * src->ne[3];
    if (nr == 0 || nc == 0) return 0;
    if ((size_t)nc * 4 > 200 * 1024) return -1;
    const unsigned n_head = (unsigned)src->ne[2];
    const unsigned n_head_log2 = 1u << (unsigned)floorf(log2f((float)n_head));
    const float m0 = powf(2.0f, -max_bias / n_head_log2), m1 = powf(2.0f, -(max_bias / 2.0f) / n_head_log2);
    int threads = 32;
    while (threads < nc && threads < 1024) threads *= 2;
    cudaStream_t st = (cudaStream_t)stream;
    const size_t smem = (size_t)nc * 4;
    if (smem > 48 * 1024) {
        static bool set[16] = {false};   // per device
        int dev = 0;
        cudaGetDevice(&dev); dev &= 15;
        if (!set[dev]) { cudaFuncSetAttribute(soft_max_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
                       cudaFuncSetAttribute(soft_max_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); set[dev] = true; }
    }
    if (mask && mask->type == T_F16)
        return rlaunch(soft_max_kernel<__half>, (unsigned)nr, threads, smem, st, (const float *)src->data, (const __half *)mask->data,
                       (float *)dst->data, nc, src->ne[1], src->ne[2], scale, max_bias, m0, m1, n_head_log2);
    return rlaunch(soft_max_kernel<float>, (unsigned)nr, threads, smem, st, (const float *)src->data, mask ? (const float *)mask->data : nullptr,
                   (float *)dst->data, nc, src->ne[1], src->ne[2], scale, max_bias, m0, m1, n_head_log2);
}

extern "C" int b200_op_argsort(const b200_tensor * src, const b200_tensor * dst, int order, void * stream) {
    if (src->type != T_F32 || dst->type != T_I32 || !contig_f(src, 4) || !contig_f(dst, 4) || src->ne[0] > 1024) return -1;
    const int ncols = (int)src->ne[0];
    const int64_t nr = src->ne[1] * src->ne[2] * src->ne[3];
    if (nr == 0 || ncols == 0) return 0;
    int npad = 1;
    while (npad < ncols) npad *= 2;
    return rlaunch(argsort_kernel, (unsigned)nr, std::max(32, npad), (size_t)npad * 4, (cudaStream_t)stream,
                   (const float *)src->data, (int32_t *)dst->data, ncols, npad, order);
}

extern "C" int b200_op_sum_rows(const b200_tensor * src, const b200_tensor * dst, void * stream) {
    if (src->type != T_F32 || dst->type != T_F32 || src->nb[0] != 4) return -1;
    const int64_t nr = src->ne[1] * src->ne[2] * src->ne[3];
    if (nr == 0) return 0;
    return rlaunch(sum_rows_kernel, (unsigned)((nr * 32 + 255) / 256), 256, 0, (cudaStream_t)stream, rmk(src), rmk(dst), nr);
}
