// quantize.cu -- stand-alone activation quantisation and weight dequantisation.
// Element-wise, HBM-bound helpers; the hot path fuses the q8_1 step into the GEMV/GEMM prologues.
#include "mmvq_units.cuh"
#include "dequant.cuh"
#include "../../include/b200_ops.h"
#include <algorithm>

namespace b200 {


// One warp per q8_1 block, exactly the arithmetic of reference ggml-cuda/quantize.cu:4-49:
// d = amax/127, q = roundf(x/d), s = xor-butterfly sum of the unquantised x, both stored as half.
__global__ void quantize_q8_1_kernel(const float * __restrict__ x, int64_t xs, uint8_t * __restrict__ y,
                                     int64_t k, int64_t k_pad, int64_t nrows) {
    const int lane = threadIdx.x & 31;
    const int64_t nb = k_pad >> 5;
    const int64_t wid = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (wid >= nb * nrows) return;
    const int64_t row = wid / nb, ib = wid - row * nb, i = ib * 32 + lane;
    const float v = i < k ? x[row * xs + i] : 0.0f;
    const float amax = warp_max(fabsf(v));
    const float sum = warp_sum(v);
    const float d = __fdiv_rn(amax, 127.0f);
    const int8_t q = amax == 0.0f ? 0 : (int8_t)roundf(__fdiv_rn(v, d));
    uint8_t * blk = y + (row * nb + ib) * 36;
    blk[4 + lane] = (uint8_t)q;
    if (lane == 0) {
        reinterpret_cast<__half *>(blk)[0] = __float2half_rn(d);
        reinterpret_cast<__half *>(blk)[1] = __float2half_rn(sum);
    }
}

// the GEMV's activation layout, written to global memory (pre-quantised path of mmvq.cuh)
__global__ void __launch_bounds__(256) quantize_acts_kernel(const float * __restrict__ x, int64_t xs, uint8_t * __restrict__ out,
                                                            int K, int M, int vec_ok) {
    pdl_trigger();
    pdl_wait();
    const uint32_t acs = (uint32_t)((K + 255) / 256) * ACT_ITEM;
    quantize_acts(x, xs, M, K >> 5, vec_ok != 0, out, acs, blockIdx.x * 256 + threadIdx.x, gridDim.x * 256);
}

template <int TYPE, typename OUT>
__global__ void dequantize_rows_kernel(const uint8_t * __restrict__ src, int64_t src_nb1, OUT * __restrict__ dst,
                                       int64_t dst_stride, int64_t k, int64_t nrows) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= k * nrows) return;
    const int64_t row = i / k, e = i - row * k;
    const float v = dequant_elem<TYPE>(src + row * src_nb1, e);
    if constexpr (sizeof(OUT) == 2) dst[row * dst_stride + e] = __float2half_rn(v);
    else dst[row * dst_stride + e] = v;
}

template <int TYPE> static int dequant_launch(const void * src, int64_t nb1, void * dst, int dst_type, int64_t ds,
                                              int64_t k, int64_t nrows, cudaStream_t st) {
    const int64_t n = k * nrows;
    const int threads = 256;
    const unsigned blocks = (unsigned)((n + threads - 1) / threads);
    if (dst_type == T_F16) dequantize_rows_kernel<TYPE, __half><<<blocks, threads, 0, st>>>((const uint8_t *)src, nb1, (__half *)dst, ds, k, nrows);
    else dequantize_rows_kernel<TYPE, float><<<blocks, threads, 0, st>>>((const uint8_t *)src, nb1, (float *)dst, ds, k, nrows);
    ++g_launches;
    return (int)cudaGetLastError();
}

}  // namespace b200

using namespace b200;

extern "C" int b200_quantize_q8_1(const float * x, int64_t xs, void * y, int64_t k, int64_t k_pad, int64_t nrows, void * stream) {
    if (k_pad % 32 || k > k_pad || nrows <= 0) return -1;
    const int64_t warps = (k_pad / 32) * nrows;
    const int threads = 256;
    const unsigned blocks = (unsigned)((warps * 32 + threads - 1) / threads);
    quantize_q8_1_kernel<<<blocks, threads, 0, (cudaStream_t)stream>>>(x, xs, (uint8_t *)y, k, k_pad, nrows);
    ++g_launches;
    return (int)cudaGetLastError();
}

extern "C" int64_t b200_quantized_acts_bytes(int K, int M) { return (int64_t)M * ((K + 255) / 256) * ACT_ITEM; }

extern "C" int b200_quantize_acts(const float * x, int64_t xs, void * out, int K, int M, void * stream) {
    if (K <= 0 || K % 32 || M <= 0) return -1;
    const int vec_ok = ((((uintptr_t)x) | ((uintptr_t)xs * 4)) & 15) == 0;
    const int groups = (K / 32) * M;                       // one 8-lane group per 32-block, 4 blocks per group
    const unsigned blocks = (unsigned)std::max(1, std::min(2 * 148, (groups / 4 + 31) / 32));
    cudaLaunchConfig_t lc = {};
    lc.gridDim = dim3(blocks); lc.blockDim = dim3(256); lc.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    lc.attrs = at; lc.numAttrs = pdl_default() ? 1 : 0;
    B200_CUDA_CHECK(cudaLaunchKernelEx(&lc, quantize_acts_kernel, x, xs, (uint8_t *)out, K, M, vec_ok));
    ++g_launches;
    return 0;
}

extern "C" int b200_dequantize_rows(int type, const void * src, int64_t src_nb1, void * dst, int dst_type, int64_t dst_stride,
                                    int64_t k, int64_t nrows, void * stream) {
    if (dst_type != T_F32 && dst_type != T_F16) return -1;
    if (k <= 0 || nrows <= 0 || k % block_elems(type)) return -1;
    cudaStream_t st = (cudaStream_t)stream;
    switch (type) {
        case T_Q4_0: return dequant_launch<T_Q4_0>(src, src_nb1, dst, dst_type, dst_stride, k, nrows, st);
        case T_Q8_0: return dequant_launch<T_Q8_0>(src, src_nb1, dst, dst_type, dst_stride, k, nrows, st);
        case T_Q4_1: return dequant_launch<T_Q4_1>(src, src_nb1, dst, dst_type, dst_stride, k, nrows, st);
        case T_Q2_K: return dequant_launch<T_Q2_K>(src, src_nb1, dst, dst_type, dst_stride, k, nrows, st);
        case T_IQ4_NL: return dequant_launch<T_IQ4_NL>(src, src_nb1, dst, dst_type, dst_stride, k, nrows, st);
        case T_IQ4_XS: return dequant_launch<T_IQ4_XS>(src, src_nb1, dst, dst_type, dst_stride, k, nrows, st);
        case T_Q3_K: return dequant_launch<T_Q3_K>(src, src_nb1, dst, dst_type, dst_stride, k, nrows, st);
        case T_Q5_0: return dequant_launch<T_Q5_0>(src, src_nb1, dst, dst_type, dst_stride, k, nrows, st);
        case T_Q5_1: return dequant_launch<T_Q5_1>(src, src_nb1, dst, dst_type, dst_stride, k, nrows, st);
        case T_Q4_K: return dequant_launch<T_Q4_K>(src, src_nb1, dst, dst_type, dst_stride, k, nrows, st);
        case T_Q5_K: return dequant_launch<T_Q5_K>(src, src_nb1, dst, dst_type, dst_stride, k, nrows, st);
        case T_Q6_K: return dequant_launch<T_Q6_K>(src, src_nb1, dst, dst_type, dst_stride, k, nrows, st);
    }
    return -1;
}
