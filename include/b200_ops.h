/*
 * b200_ops.h -- plain C ABI of the B200 (sm_100a) kernels for llamacog's quantized forward pass.
 *
 * libb200ops.so exports these; no ggml, torch or C++ types cross the boundary: device pointers,
 * sizes, ggml type ids as ints (reference ggml/include/ggml.h:352-366) and a CUDA stream handle
 * passed as void*.  All functions are asynchronous on `stream` and return 0 on success, a negative
 * value for an unsupported argument combination, or a positive cudaError_t.
 *
 * The ggml backend plugin (include/ggml-b200.h, libggml-b200.so) is a thin dispatcher from
 * ggml_tensor nodes onto these entry points; tests and bench.py bind them with ctypes.
 * Each entry cites the reference interface it replaces (paths relative to /root/reference).
 */
#ifndef B200_OPS_H
#define B200_OPS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_TYPE_F32   0
#define B200_TYPE_F16   1
#define B200_TYPE_Q4_0  2
#define B200_TYPE_Q8_0  8
#define B200_TYPE_Q8_1  9
#define B200_TYPE_Q4_K 12
#define B200_TYPE_Q5_K 13
#define B200_TYPE_Q6_K 14
#define B200_TYPE_I32  26

/* ---- MUL_MAT, decode regime (activation columns M small) ----------------------------------------
 * Replaces ggml_cuda_mul_mat_vec_q + quantize_row_q8_1_cuda (ggml/src/ggml-cuda/mmvq.cu:491-561,
 * quantize.cu:145-157) and, with ids, the single-token MUL_MAT_ID path (mmvq.cu:161-164,547-554).
 *   dst[ch][m*ds + n] = sum_k dequant(W[wsel(ch)][n][k]) * q8_1(X[ch % x_ch_mod][m])[k]
 * W rows are GGUF blocks of `type`, row n at W + wsel*w_ch_stride + n*nb1 (bytes, 2-byte aligned);
 * X is fp32, column m at X + xch*x_ch_stride + m*xs (floats); the q8_1 quantisation of X is fused
 * into the kernel prologue.  M > 8 is processed in several passes over W. */
typedef struct b200_mmvq_args {
    int type;
    const void * W;
    int64_t nb1;            /* bytes between weight rows */
    const float * X;
    int64_t xs;             /* floats between activation columns */
    float * dst;
    int64_t ds;             /* floats between dst columns */
    int N, K, M;
    int nchannels;          /* >= 1: batch / expert-slot channels (grid.y) */
    int64_t w_ch_stride;    /* bytes between weight matrices */
    int64_t x_ch_stride;    /* floats */
    int64_t d_ch_stride;    /* floats */
    const int32_t * ids;    /* optional device array [nchannels]: weight matrix per channel */
    int x_ch_mod;           /* channel c reads X channel c % x_ch_mod (0: no wrap) */
    int w_ch_div;           /* without ids: channel c reads weight matrix c / w_ch_div (0: 1) */
    int pdl;                /* 1: launch with programmatic dependent launch, 0: off, -1: env B200_PDL */
    int32_t * dbg_sumi;     /* optional device [M][N][K/32] zero-initialised: receives the exact int32
                               block dot products (reference vecdotq.cuh sumi/dot1) for parity tests */
    const void * Xq;        /* optional: activations already quantised by b200_quantize_acts (then X may be
                               NULL): [channels][M][ceil(K/256)] items of 400 bytes; the kernel then copies
                               them instead of quantising X itself (better when K is large and N small) */
    int64_t xq_ch_stride;   /* bytes between channels of Xq */
} b200_mmvq_args;

int b200_mul_mat_vec_q_ex(const b200_mmvq_args * args, void * stream);
int b200_mul_mat_vec_q(int type, const void * W, int64_t nb1, const float * X, int64_t xs,
                       float * dst, int64_t ds, int N, int K, int M, void * stream);
int b200_mmvq_supported_type(int type);

/* ---- activation quantisation as a stand-alone op ---------------------------------------------
 * Replaces quantize_row_q8_1_cuda (ggml/src/ggml-cuda/quantize.cu:4-49,145-157): x fp32 [nrows][k]
 * (row stride xs floats) -> block_q8_1 rows of k_pad/32 blocks, zero padded beyond k. */
int b200_quantize_q8_1(const float * x, int64_t xs, void * y, int64_t k, int64_t k_pad, int64_t nrows, void * stream);

/* The same quantisation in the shared-memory layout of the GEMV (per 32 values: 32 int8, fp32 d,
 * d*sum(q), -32*sum(q[0..15]), -32*sum(q[16..31]); 400 bytes per 256 values), for b200_mmvq_args.Xq.
 * x fp32 [M][K] (column stride xs floats) -> out [M][ceil(K/256)*400 bytes]. */
int b200_quantize_acts(const float * x, int64_t xs, void * out, int K, int M, void * stream);
int64_t b200_quantized_acts_bytes(int K, int M);

/* ---- dequantisation ---------------------------------------------------------------------------
 * Replaces dequantize_row_<type> (ggml/src/ggml-quants.c:249,343,1274,1480,1684; GPU analogue
 * ggml-cuda/convert.cu): nrows rows of k elements, row r at src + r*src_nb1 bytes, output fp32 (or
 * fp16 when dst_type == B200_TYPE_F16) row r at dst + r*dst_stride elements.  Bit-exact to the
 * reference for fp32 output. */
int b200_dequantize_rows(int type, const void * src, int64_t src_nb1, void * dst, int dst_type, int64_t dst_stride,
                         int64_t k, int64_t nrows, void * stream);

/* number of kernels launched by this library so far in this process (bench.py gpu_launches) */
int b200_kernel_launch_count(void);

#ifdef __cplusplus
}
#endif
#endif /* B200_OPS_H */
