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
__device__ __forceinline__ void soft_max_row(const float * __restrict__ x, const MT * __restrict__ mask, float * __restrict__ y,
                                             int nc, int64_t ne01, int64_t ne02, float scale, float max_bias, float m0, float m1,
                                             unsigned n_head_log2, float * row, float * red) {
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
template <typename MT>
__global__ void __launch_bounds__(1024) soft_max_kernel(const float * __restrict__ x, const MT * __restrict__ mask, float * __restrict__ y,
                                                        int nc, int64_t ne01, int64_t ne02, float scale, float max_bias, float m0, float m1,
                                                        unsigned n_head_log2) {
    pdl_trigger(); pdl_wait();
    extern __shared__ float row[];
    __shared__ float red[32];
    soft_max_row<MT>(x, mask, y, nc, ne01, ne02, scale, max_bias, m0, m1, n_head_log2, row, red);
}

// ---- argsort (reference ggml-cpu/ops.cpp ggml_compute_forward_argsort_f32): bitonic sort of one row per CTA ----
// (no __restrict__ here: the fused kernel below sorts values that the same launch has just written)
__device__ __forceinline__ void argsort_row(const float * x, int32_t * dst, int ncols, int npad, int desc, int * idx) {
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
__global__ void __launch_bounds__(1024) argsort_kernel(const float * __restrict__ x, int32_t * __restrict__ dst, int ncols, int npad, int desc) {
    pdl_trigger(); pdl_wait();
    extern __shared__ int idx[];
    argsort_row(x, dst, ncols, npad, desc, idx);
}
// SOFT_MAX (no mask) of a row followed by the ARGSORT of the result -- the router of a MoE layer (reference src/llama-graph.cpp:663-693:
// ggml_soft_max, ggml_top_k) -- in one launch: the same two bodies one after the other, so both tensors get the bits of the two kernels.
__global__ void __launch_bounds__(1024) soft_max_argsort_kernel(const float * __restrict__ x, float * y, int32_t * dst, int nc, int npad,
                                                                int64_t ne01, int64_t ne02, float scale, int desc) {
    pdl_trigger(); pdl_wait();
    extern __shared__ float row[];
    __shared__ float red[32];
    soft_max_row<float>(x, nullptr, y, nc, ne01, ne02, scale, 0.0f, 1.0f, 1.0f, 1u, row, red);
    __syncthreads();                     // y of this row is complete (written by this CTA) and the shared row is free again
    argsort_row(y, dst, nc, npad, desc, reinterpret_cast<int *>(row));
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

// ---- MoE decode: the weights of the chosen experts and the weighted sum of their outputs in ONE launch ----
// Replaces GET_ROWS(probs, ids) -> SUM_ROWS -> DIV -> MUL(experts, weights) -> ADD x (n_used - 1) of build_moe_ffn (reference
// src/llama-graph.cpp:696-775; five launches of 3-4 us each per layer and token).  Every tensor of those nodes is written with the
// bits the single kernels give it: the copy of get_rows_kernel<F32>, the lane-strided sum + butterfly of sum_rows_kernel, the IEEE
// division / product / sum of binary_kernel (explicit roundings: nothing may contract to an FMA here).
struct moe_comb {
    const char * probs; int64_t probs_nb2;       // probs as [1, n_expert, T]: element e of token t at probs + 4 e + t probs_nb2
    const char * ids; int64_t ids_nb0, ids_nb1;  // chosen experts [n_used, T] (a view of the argsort result)
    const float * experts;                       // [n_embd, n_used, T] contiguous
    float * w; float * wsum; float * wnorm;      // GET_ROWS dst [n_used, T], SUM_ROWS dst [T], DIV dst [n_used, T], contiguous
    float * weighted;                            // MUL dst [n_embd, n_used, T] contiguous (may be `experts`)
    float * adds[7];                             // ADD dsts [n_embd, T]: adds[k] = (k ? adds[k - 1] : weighted[:, 0]) + weighted[:, k + 1]
    int n_embd, n_used, T;
};
// One CTA per token: the chosen experts and their probabilities are read into registers by every warp BEFORE the barrier, everything
// is written after it -- so for one token the outputs may lie anywhere over `ids` / `probs` (dead after GET_ROWS in the graph's own
// schedule: the allocator hands their memory to SUM_ROWS / DIV at once).
__global__ void __launch_bounds__(1024) moe_combine_kernel(const moe_comb p) {
    pdl_trigger(); pdl_wait();
    const int t = blockIdx.y, lane = threadIdx.x & 31;
    float wv = 0.0f;
    if (lane < p.n_used) {
        const int32_t e = *(const int32_t *)(p.ids + lane * p.ids_nb0 + t * p.ids_nb1);
        wv = *(const float *)(p.probs + (int64_t)e * 4 + t * p.probs_nb2);
    }
    float s = 0.0f;
    if (lane < p.n_used) s += wv;                  // sum_rows_kernel: lane i starts from 0 and adds x[i], then the butterfly
    s = warp_sum(s);
    const float wn = lane < p.n_used ? __fdiv_rn(wv, s) : 0.0f;
    float wk[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) wk[k] = __shfl_sync(0xffffffffu, wn, k);
    __syncthreads();
    if (threadIdx.x < 32) {
        if (lane < p.n_used) { p.w[t * p.n_used + lane] = wv; p.wnorm[t * p.n_used + lane] = wn; }
        if (lane == 0) p.wsum[t] = s;
    }
    for (int e = threadIdx.x; e < p.n_embd; e += blockDim.x) {
        float m[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) m[k] = __fmul_rn(k < p.n_used ? p.experts[((int64_t)t * p.n_used + k) * p.n_embd + e] : 0.0f, wk[k]);
        // all reads of element e are done: the stores follow the node order, so outputs the graph allocator laid over one another
        // (MUL in place on the experts, an ADD in place on a slot or on the running sum) end up as after the separate launches
#pragma unroll
        for (int k = 0; k < 8; ++k) if (k < p.n_used) p.weighted[((int64_t)t * p.n_used + k) * p.n_embd + e] = m[k];
        float acc = m[0];
#pragma unroll
        for (int k = 1; k < 8; ++k) if (k < p.n_used) { acc = __fadd_rn(acc, m[k]); p.adds[k - 1][(int64_t)t * p.n_embd + e] = acc; }
    }
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
        B200_GR_CASE(T_Q4_K) B200_GR_CASE(T_Q5_K) B200_GR_CASE(T_Q6_K) B200_GR_CASE(T_Q4_1) B200_GR_CASE(T_Q5_0) B200_GR_CASE(T_Q5_1) B200_GR_CASE(T_Q2_K) B200_GR_CASE(T_Q3_K) B200_GR_CASE(T_IQ4_NL) B200_GR_CASE(T_IQ4_XS)
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

extern "C" int b200_op_moe_combine(const b200_tensor * probs, const b200_tensor * ids, const b200_tensor * experts, float * w, float * wsum, float * wnorm,
                                   float * weighted, float * const * adds, void * stream) {
    const int n_embd = (int)experts->ne[0], n_used = (int)experts->ne[1], T = (int)experts->ne[2];
    if (probs->type != T_F32 || ids->type != T_I32 || experts->type != T_F32 || experts->ne[3] != 1 || !contig_f(experts, 4)) return -1;
    if (n_used < 2 || n_used > 8 || T < 1 || T > 65535 || n_embd < 1 || ids->ne[0] != n_used || ids->ne[1] != T || probs->ne[0] != 1 || probs->ne[2] != T) return -1;
    if (probs->nb[1] != 4 || !w || !wsum || !wnorm || !weighted || !adds) return -1;
    moe_comb p = {};
    p.probs = (const char *)probs->data; p.probs_nb2 = probs->nb[2];
    p.ids = (const char *)ids->data; p.ids_nb0 = ids->nb[0]; p.ids_nb1 = ids->nb[1];
    p.experts = (const float *)experts->data; p.w = w; p.wsum = wsum; p.wnorm = wnorm; p.weighted = weighted;
    for (int k = 0; k + 1 < n_used; ++k) { if (!adds[k]) return -1; p.adds[k] = adds[k]; }
    p.n_embd = n_embd; p.n_used = n_used; p.T = T;
    const int threads = n_embd >= 4096 ? 1024 : n_embd >= 1024 ? 512 : 256;
    return rlaunch(moe_combine_kernel, dim3(1, (unsigned)T), threads, 0, (cudaStream_t)stream, p);
}

extern "C" int b200_op_soft_max_argsort(const b200_tensor * src, const b200_tensor * probs, const b200_tensor * order, float scale, int desc, void * stream) {
    if (src->type != T_F32 || probs->type != T_F32 || order->type != T_I32 || !contig_f(src, 4) || !contig_f(probs, 4) || !contig_f(order, 4)) return -1;
    const int nc = (int)src->ne[0];
    const int64_t nr = src->ne[1] * src->ne[2] * src->ne[3];
    if (nc < 1 || nc > 1024 || nr < 1 || probs->ne[0] != nc || order->ne[0] != nc) return -1;
    int npad = 1;
    while (npad < nc) npad *= 2;
    const int threads = std::max(32, npad);          // what b200_op_soft_max and b200_op_argsort launch for this width
    return rlaunch(soft_max_argsort_kernel, (unsigned)nr, threads, (size_t)npad * 4, (cudaStream_t)stream, (const float *)src->data, (float *)probs->data,
                   (int32_t *)order->data, nc, npad, src->ne[1], src->ne[2], scale, desc);
}
