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
#define B200_TYPE_Q4_1  3
#define B200_TYPE_Q5_0  6
#define B200_TYPE_Q5_1  7
#define B200_TYPE_Q8_0  8
#define B200_TYPE_Q8_1  9
#define B200_TYPE_Q2_K 10
#define B200_TYPE_Q3_K 11
#define B200_TYPE_Q4_K 12
#define B200_TYPE_Q5_K 13
#define B200_TYPE_Q6_K 14
#define B200_TYPE_IQ4_NL 20
#define B200_TYPE_IQ4_XS 23
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
    /* batch / expert-slot channels: a (ny x nz) grid of independent GEMVs.  Channel (y, z) uses
     *   weight matrix  W + ids[z*ids_sz + y] * w_sy                    when ids != NULL (MUL_MAT_ID), else
     *                  W + (y / w_div_y) * w_sy + (z / w_div_z) * w_sz   (ne02/ne03 broadcast of MUL_MAT)
     *   activations    X + (y % x_mod_y) * x_sy + z * x_sz  (floats)    (ne11 broadcast of MUL_MAT_ID)
     *   output         dst + y * d_sy + z * d_sz  (floats) */
    int ny, nz;             /* 0 is treated as 1 */
    int64_t w_sy, w_sz;     /* bytes */
    int w_div_y, w_div_z;   /* 0 is treated as 1 */
    int64_t x_sy, x_sz;     /* floats */
    int x_mod_y;            /* 0: no wrap */
    int64_t d_sy, d_sz;     /* floats */
    const int32_t * ids;    /* optional device array */
    int64_t ids_sz;         /* ints between z rows of ids */
    int pdl;                /* 1: launch with programmatic dependent launch, 0: off, -1: env B200_PDL (default on) */
    int32_t * dbg_sumi;     /* optional device [M][N][K/32]: receives the exact int32 block dot products
                               (reference vecdotq.cuh sumi/dot1) for parity tests; single channel only */
    const void * Xq;        /* optional: activations already quantised by b200_quantize_acts (then X may be
                               NULL): per channel [M][ceil(K/256)] items of 400 bytes; the kernel then copies
                               them instead of quantising X itself (better when K is large and N small) */
    int64_t xq_sy, xq_sz;   /* bytes between channels of Xq (same y wrap as X) */
    /* optional hint: byte ranges (the next matmul's weights) that the CTAs pull into L2 with cp.async.bulk.prefetch
     * once they have issued their own last weight load, so that HBM keeps streaming through the kernel's tail and the
     * small kernels that follow.  No effect on results. */
    int pf_n;               /* 0..3 */
    const void * pf_ptr[3];
    int64_t pf_bytes[3];
    /* Decode-step fusions of the batch-1 kernel (mmvq_r.cuh); honoured only when b200_mmvq_fusable() says so for the same
     * arguments, otherwise the call fails with -4.  A kernel hop of a decode step costs 3.5-5 us whatever it computes
     * (profiles/r01_ktrace_tg_v5.txt), so the element-wise neighbours of a mat-vec ride along with it:
     *  - epilogue ADD (single matrix, M = 1): fuse_sum[n] = dst[n] + fuse_resid[n]; dst is still written.
     *    Replaces the GGML_OP_ADD that follows wo / ffn_down (reference ggml-cuda/binbcast.cu).
     *  - prologue RMS_NORM + MUL (M = 1, K <= 4096): the activations are rms_norm(norm_x, norm_eps) * norm_w; every CTA
     *    computes them for itself, CTA 0 also stores the normalised row (norm_y), the product (norm_out) and its
     *    quantised form into Xq (which is then an OUTPUT of the call, for later mat-vecs on the same column).  Same
     *    arithmetic, bit for bit, as b200_op_add_rms_norm_mul without the ADD.  norm_x must not alias the outputs. */
    const float * fuse_resid; float * fuse_sum;
    const float * norm_x; const float * norm_w; float * norm_y; float * norm_out; float norm_eps;
    /*  - epilogue ROPE + KV-cache store of a q/k/v launch (M = 1, see struct b200_gemv_rope below) */
    const struct b200_gemv_rope * rope;
    /* Row-split hand-over folded into the launch (all optional; protocol of csrc/peer.cu, flags are 32-bit words, zero when idle,
     * in memory the launching device can reach).  Replaces the wait / signal launches around a shard's mat-vec, i.e. the
     * cudaStreamWaitEvent / cudaEventRecord pairs of the reference's split MUL_MAT (ggml-cuda/ggml-cuda.cu:1628-1630, :1724-1745).
     * Needs pre-quantised activations (Xq).
     *   hs_wait    every CTA spins until *hs_wait != 0 before it reads Xq            (peer: "the activations have arrived")
     *   hs_counter zeroed word in the launching device's memory; elects the CTA that finishes last, which then, in this order,
     *   hs_join    waits for hs_join[t * hs_join_stride], t in hs_join_mask, and lowers them  (main: "every peer's slice is written")
     *   hs_clear   stores 0 to *hs_clear                                             (peer: lowers the wait flag)
     *   hs_signal  system-scope fence, then *hs_signal = 1                           (peer: "my slice of dst is written")
     *   hs_timeout_ms / hs_error: spin budget (0 = none); on expiry *hs_error = 1 and the kernel carries on */
    const uint32_t * hs_wait; uint32_t * hs_clear; uint32_t * hs_signal;
    uint32_t * hs_join; int hs_join_stride; uint32_t hs_join_mask;
    unsigned * hs_counter; unsigned hs_timeout_ms; uint32_t * hs_error;
    /* gate / up of a decode step as a group launch (b200_mul_mat_vec_q_group, ng = 2: matrix 0 = gate, 1 = up, equal N, N % 32 == 0,
     * batch-1 kernel only, see b200_mmvq_fusable): every CTA takes the same whole 32-row blocks of BOTH matrices and also computes
     *   sm_act[n] = unary(gate[n]),  sm_out[n] = sm_act[n] * up[n],  and (sm_xq != NULL) the q8 activations of sm_out
     * with the arithmetic of b200_op_unary_mul (bit-identical); replaces that launch (reference: unary.cu silu + binbcast.cu mul).
     * sm_act == NULL: off.  sm_op: B200_UNARY_SILU only.  sm_act / sm_out may alias the gate / up results.
     * sm_counters: unused (the first version counted blocks across CTAs there; kept so that the struct layout stays). */
    int sm_op; float * sm_act; float * sm_out; void * sm_xq; unsigned * sm_counters;
    /* the same for the chosen experts of ONE token of a MoE layer (b200_mul_mat_vec_q_ex with ids, ny = n_used <= 3; reference
     * ggml-cuda.cu:2007-2148 mul_mat_id x 2 + unary.cu + binbcast.cu): W = the gate stack, sm_up_W = the up stack (same geometry),
     * dst / sm_up_dst = the gate / up results [N, ny] (slot stride d_sy), sm_act / sm_out likewise, sm_xq = ny q8 columns of N / 256
     * items (N % 256 == 0).  One launch instead of three. */
    const void * sm_up_W; float * sm_up_dst;
} b200_mmvq_args;

/* 1 when b200_mul_mat_vec_q_ex (ng = 1, W/N NULL) or b200_mul_mat_vec_q_group would run these arguments on the
 * register-resident batch-1 kernel, i.e. when the fuse_* / norm_* fields may be used.  Host-side check only. */
int b200_mmvq_fusable(const b200_mmvq_args * args, int ng, const void * const * W, const int * N);

int b200_mul_mat_vec_q_ex(const b200_mmvq_args * args, void * stream);
int b200_mul_mat_vec_q(int type, const void * W, int64_t nb1, const float * X, int64_t xs,
                       float * dst, int64_t ds, int N, int K, int M, void * stream);
/* ng <= 3 weight matrices of one type, K and row stride times the same activations (args: type, nb1, X / Xq, xs, K, M <= 8, pdl)
 * in ONE launch: matrix g has N[g] rows at W[g] and writes dst[g][m*ds[g] + n].  The q/k/v and gate/up projections of a decode
 * step (the reference launches them separately: one mul_mat_vec_q per MUL_MAT node, ggml-cuda/mmvq.cu:491-561). */
int b200_mul_mat_vec_q_group(const b200_mmvq_args * args, int ng, const void * const * W, float * const * dst, const int * N,
                             const int64_t * ds, void * stream);
int b200_mmvq_supported_type(int type);

/* ---- kernel timeline trace (debug tooling, not on any product path) ----------------------------
 * After b200_ktrace_enable() thread 0 of block 0 of every kernel of this library records the global timer at kernel
 * entry and when it passes its programmatic-dependent-launch wait; b200_ktrace_dump writes the events as text
 * ("<ns> <gridid> <event> <ctas> <threads>" per line).  Shows a decode step as it runs under CUDA-graph replay with PDL
 * overlap, which a serialising profiler cannot.  The plugin enables it when B200_KTRACE=<file> is set. */
int b200_ktrace_enable(void);
int b200_ktrace_dump(const char * path);


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


/* ==== tensor-level operators (the non-matmul nodes of the Llama / Mixtral graphs and the MUL_MAT
 * dispatch).  A b200_tensor mirrors the fields of ggml_tensor that an op needs (reference
 * ggml/include/ggml.h:569-604): device data pointer, ggml type id, ne[4] elements and nb[4] byte
 * strides.  Every function is asynchronous on `stream`, returns 0 on success, -1 when the argument
 * combination is not supported (the plugin's supports_op asks the matching *_supported predicate
 * first) or a positive cudaError_t. ============================================================ */
typedef struct b200_tensor {
    void *  data;
    int     type;
    int64_t ne[4];
    int64_t nb[4];
} b200_tensor;

#define B200_TYPE_BF16 30

enum b200_binary_op { B200_ADD = 0, B200_SUB = 1, B200_MUL = 2, B200_DIV = 3 };
/* unary op ids = ggml_unary_op (reference ggml/include/ggml.h:525-543) */
enum b200_unary_op { B200_UNARY_ABS = 0, B200_UNARY_SGN, B200_UNARY_NEG, B200_UNARY_STEP, B200_UNARY_TANH, B200_UNARY_ELU,
                     B200_UNARY_RELU, B200_UNARY_SIGMOID, B200_UNARY_GELU, B200_UNARY_GELU_QUICK, B200_UNARY_SILU,
                     B200_UNARY_HARDSWISH, B200_UNARY_HARDSIGMOID, B200_UNARY_EXP, B200_UNARY_GELU_ERF };

/* dst = a (op) b with b broadcast over a; f32.  Replaces ggml_cuda_op_add/sub/mul/div (ggml-cuda/binbcast.cu:26-). */
int b200_op_binary(int op, const b200_tensor * a, const b200_tensor * b, const b200_tensor * dst, void * stream);
/* element-wise f32 on contiguous tensors.  Replaces ggml-cuda/unary.cu:87-. */
int b200_op_unary(int op, const b200_tensor * src, const b200_tensor * dst, void * stream);
/* dst = src * scale.  Replaces ggml-cuda/scale.cu:3-. */
int b200_op_scale(const b200_tensor * src, const b200_tensor * dst, float scale, void * stream);
/* row-wise x / sqrt(mean(x^2) + eps).  Replaces ggml_cuda_op_rms_norm (ggml-cuda/norm.cu:108-). */
int b200_op_rms_norm(const b200_tensor * src, const b200_tensor * dst, float eps, void * stream);
/* rotary embedding, modes normal (0) and neox (2), f32 or f16 data, optional frequency factors.
 * Replaces ggml_cuda_op_rope (ggml-cuda/rope.cu:41,85-); parameters as ggml_rope_ext packs them
 * into op_params (reference ggml/src/ggml-cpu/ops.cpp:5187-5202). */
typedef struct b200_rope_params {
    int n_dims, mode, n_ctx_orig;
    float freq_base, freq_scale, ext_factor, attn_factor, beta_fast, beta_slow;
} b200_rope_params;
int b200_op_rope(const b200_tensor * src, const b200_tensor * pos, const b200_tensor * freq_factors, const b200_tensor * dst,
                 const b200_rope_params * prm, void * stream);
/* ROPE and the KV-cache stores of a decode step inside the q/k/v mat-vec (b200_mmvq_args.rope; batch-1 kernel only, checked by
 * b200_mmvq_fusable).  Per matrix g of the launch, with n the row index (= flat element index of the [head_dim, n_head] result):
 *   mode 0: nothing            mode 1: out[n] = rope(dst)[n]            mode 2: mode 1 and cell[n] = convert(out[n])  (K row)
 *   mode 3: cell[n] = convert(dst[n])  (V row)
 * rope in mode "normal" (adjacent pairs) at position pos[0]; cell_type f32 / f16 / bf16, contiguous cell; cell_slot (optional):
 * device address of a pointer that replaces cell when the kernel runs (CUDA-graph replay, see b200_op_cpy_indirect).  The raw
 * dst of a roped matrix is not written.  Bit-identical to b200_op_rope_cpy_batch on the mat-vec's output; replaces the
 * reference's rope.cu + cpy.cu launches that follow the q/k/v projections. */
typedef struct b200_gemv_rope {
    int mode[3];
    float * out[3];
    void * cell[3]; void * const * cell_slot[3]; int cell_type[3];
    int head_dim[3];
    const int32_t * pos; const float * ffac;
    b200_rope_params prm;
} b200_gemv_rope;
/* copy with type conversion and arbitrary strides (CPY / CONT / DUP): f32,f16,bf16 -> f32,f16,bf16; f32 -> q8_0 (the
 * KV-cache store; bit-exact to quantize_row_q8_0_ref, ggml-quants.c:191-212); same-type block copies of quantised
 * data.  Replaces ggml_cuda_cpy (ggml-cuda/cpy.cu:303-). */
int b200_op_cpy(const b200_tensor * src, const b200_tensor * dst, void * stream);
/* same, with the destination base address read on the device from *dst_slot when the kernel runs (dst->data is ignored):
 * lets a captured CUDA graph of a decode step be replayed while the KV-cache cell it writes moves (the reference's
 * ggml_cuda_cpy_dest_ptrs_copy indirection, ggml-cuda/cpy.cu:360-372,591-612, ggml-cuda.cu:2502-2564). */
int b200_op_cpy_indirect(const b200_tensor * src, const b200_tensor * dst, void * const * dst_slot, void * stream);
int b200_op_cpy_supported(int src_type, int dst_type);
/* dst[:, i] = dequant(src[:, ids[i]]), dst f32.  Replaces ggml_cuda_op_get_rows (ggml-cuda/getrows.cu:42-). */
int b200_op_get_rows(const b200_tensor * src, const b200_tensor * ids, const b200_tensor * dst, void * stream);
/* soft_max(src * scale + slope * mask) over rows; mask f16 or f32 or NULL.  Replaces ggml-cuda/softmax.cu:23-. */
int b200_op_soft_max(const b200_tensor * src, const b200_tensor * mask, const b200_tensor * dst, float scale, float max_bias,
                     void * stream);
/* row-wise argsort (order 0 ascending, 1 descending), dst i32; ne0 <= 1024.  Replaces ggml-cuda/argsort.cu:11-. */
int b200_op_argsort(const b200_tensor * src, const b200_tensor * dst, int order, void * stream);
/* probs = SOFT_MAX(src * scale) (no mask), order = ARGSORT(probs) in one launch: the router of a MoE layer (reference
 * src/llama-graph.cpp:663-693; ggml-cuda softmax.cu + argsort.cu).  Rows of <= 1024 contiguous f32; both outputs bit-identical to the
 * two operators. */
int b200_op_soft_max_argsort(const b200_tensor * src, const b200_tensor * probs, const b200_tensor * order, float scale, int desc, void * stream);
/* dst[0, r] = sum over row r.  Replaces ggml-cuda/sumrows.cu. */
int b200_op_sum_rows(const b200_tensor * src, const b200_tensor * dst, void * stream);
/* The tail of a MoE layer for a few tokens (reference src/llama-graph.cpp:696-775 build_moe_ffn; run there as ggml-cuda getrows.cu,
 * sumrows.cu, binbcast.cu x (2 + n_used - 1)) in one launch:
 *   w = GET_ROWS(probs [1, n_expert, T], ids [n_used, T]); wsum = SUM_ROWS(w); wnorm = w / wsum;
 *   weighted = experts [n_embd, n_used, T] * wnorm; adds[k] = (k ? adds[k - 1] : weighted[:, 0]) + weighted[:, k + 1]
 * Every output holds the bits the single operators give it.  2 <= n_used <= 8; w / wnorm [n_used, T], wsum [T], weighted like
 * experts (may be the same memory), adds[0 .. n_used - 2] [n_embd, T], all contiguous f32. */
int b200_op_moe_combine(const b200_tensor * probs, const b200_tensor * ids, const b200_tensor * experts, float * w, float * wsum, float * wnorm,
                        float * weighted, float * const * adds, void * stream);
/* softmax(scale * Q K^T + slope * mask [softcap]) V with online softmax and GQA broadcast; Q f32 [D, n_batch, n_head],
 * K/V f16 or q8_0 [D, n_kv, n_head_kv] (strided views of the KV cache), mask f16 [n_kv, >= n_batch], dst f32
 * [D, n_head, n_batch].  Replaces ggml_cuda_flash_attn_ext (ggml-cuda/fattn.cu:271-346); op contract
 * ggml/include/ggml.h:1848-1870.  scratch: device memory of b200_flash_attn_scratch_bytes() for the split-KV partials. */
int b200_op_flash_attn_ext(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask,
                           const b200_tensor * dst, float scale, float max_bias, float logit_softcap,
                           void * scratch, int64_t scratch_bytes, void * stream);
/* same, and (xq != NULL) the result also as the pre-quantised activations of the GEMV that consumes it (b200_quantize_acts
 * layout for K = D * n_head, row m at xq + m * ceil(K / 256) * 400): saves the quantisation launch of a decode step.
 * Only for shapes b200_flash_attn_xq_supported accepts (decode rows, D = 128, one sequence). */
int b200_flash_attn_xq_supported(const b200_tensor * q, const b200_tensor * k);
int b200_op_flash_attn_ext_xq(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask,
                              const b200_tensor * dst, float scale, float max_bias, float logit_softcap,
                              void * scratch, int64_t scratch_bytes, void * xq, void * stream);
/* prefill: same, and the result also as the bf16 activation rows [rows][D * n_head] of the MMQ that consumes it (the output
 * projection): saves the activation conversion launch.  -1 when the tensor-core kernel does not take the shape. */
int b200_flash_attn_xb_supported(const b200_tensor * q, const b200_tensor * k, const b200_tensor * dst);
int b200_op_flash_attn_ext_xb(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask,
                              const b200_tensor * dst, float scale, float max_bias, float logit_softcap,
                              void * scratch, int64_t scratch_bytes, void * xb, void * stream);
int64_t b200_flash_attn_scratch_bytes(const b200_tensor * q, const b200_tensor * k);
int b200_flash_attn_supported(int k_type, int v_type, int64_t dk, int64_t dv);
/* MUL_MAT: dst[N, M, b2, b3] = src0[K, N, a2, a3] x src1[K, M, b2, b3], src0 broadcast over dims 2,3.
 * Quantised src0 (Q4_0, Q4_1, Q5_0, Q5_1, Q8_0, Q2_K, Q3_K, Q4_K, Q5_K, Q6_K, IQ4_NL, IQ4_XS: b200_mmvq_supported_type) with f32 src1
 * goes to the MMVQ GEMV for M <= 8 and to the MMQ GEMM (b200_mul_mat_q_supported: K % 256 == 0; K % 2048 == 0 for Q6_K, Q2_K, Q3_K, IQ4_XS)
 * beyond; f16/bf16/f32 src0 to a plain tiled kernel.  Replaces ggml_cuda_mul_mat (ggml-cuda/ggml-cuda.cu:1936-2005).
 * scratch: device memory for the quantised activations (b200_mul_mat_scratch_bytes). */
int b200_op_mul_mat(const b200_tensor * src0, const b200_tensor * src1, const b200_tensor * dst,
                    void * scratch, int64_t scratch_bytes, void * stream);
int64_t b200_mul_mat_scratch_bytes(const b200_tensor * src0, const b200_tensor * src1);
int b200_mul_mat_supported(const b200_tensor * src0, const b200_tensor * src1);
/* MMQ: dst[m*ds + n] = sum_k dequant(W[n][k]) * X[m*xs + k] for M > 8 tokens on the tcgen05 tensor cores (weights dequantised
 * to bf16 in shared memory, activations converted to bf16 once into `scratch`, f32 accumulation in TMEM).
 * Replaces ggml_cuda_mul_mat_q / mul_mat_q<type> (ggml/src/ggml-cuda/mmq.cu:1-140, mmq.cuh:3390-3760) for the dense
 * (non-MoE) case.  Requires K % 256 == 0, W and nb1 16-byte aligned (b200_mul_mat_q_supported); scratch of
 * b200_mul_mat_q_scratch_bytes(K, M) bytes.  Weight rows may be read up to 30 bytes past a row's end (never past
 * row N-1's stride), which ggml buffers pad for (get_alloc_size). */
int b200_mul_mat_q(int type, const void * W, int64_t nb1, const float * X, int64_t xs, float * dst, int64_t ds,
                   int N, int K, int M, void * scratch, int64_t scratch_bytes, void * stream);
int64_t b200_mul_mat_q_scratch_bytes(int K, int M);
int b200_mul_mat_q_supported(int type, const void * W, int64_t nb1, int K);
/* The two halves of b200_mul_mat_q, for callers that multiply several weight matrices by the same activations
 * (q/k/v, gate/up): convert once (out: b200_acts_bf16_bytes(K, M) bytes, 128-byte aligned), then one
 * b200_mul_mat_q_bf16 per weight matrix.  The reference quantises src1 once per MUL_MAT (ggml-cuda/mmq.cu:118-140). */
int64_t b200_acts_bf16_bytes(int K, int M);
int b200_convert_acts_bf16(const float * X, int64_t xs, void * out, int K, int M, void * stream);
int b200_mul_mat_q_bf16(int type, const void * W, int64_t nb1, const void * Xb, float * dst, int64_t ds,
                        int N, int K, int M, void * stream);
/* Up to three weight matrices of one type, K and row stride times the same bf16 activations in ONE grid (q/k/v of a prefill
 * batch: the k / v projections alone are 8 row tiles, a fraction of the SMs).  dst[g][m * ds[g] + n], N[g] rows each.
 * No reference counterpart: ggml-cuda launches one mul_mat_q per MUL_MAT node (ggml-cuda.cu:1936-2005). */
int b200_mul_mat_q_bf16_group(int type, int ng, const void * const * W, int64_t nb1, const void * Xb, float * const * dst,
                              const int64_t * ds, const int * N, int K, int M, void * stream);
/* b200_mul_mat_q for callers that cannot allocate device memory (CUDA-graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  No counterpart in the reference.  Same arguments and
 * results (bit for bit) as b200_mul_mat_q without the scratch: the bf16 activation rows go to a static __device__ array of
 * B200_NOALLOC_SCRATCH_BYTES inside the library (each GPU that loads the library holds it until exit), so the call needs
 * K * M * 2 <= B200_NOALLOC_SCRATCH_BYTES and returns B200_ERR_NOALLOC_TOO_LARGE (with a line on stderr) beyond that.
 * The call itself performs no cudaMalloc / cudaFree.  All calls on a device share the array: calls that overlap on the
 * GPU (two streams, concurrent graph replays) race -- serialise them.  Nothing in the array needs clearing between calls. */
#define B200_NOALLOC_SCRATCH_BYTES (8 << 20)
#define B200_ERR_NOALLOC_TOO_LARGE (-7)
int b200_mul_mat_q_noalloc(int type, const void * W, int64_t nb1, const float * X, int64_t xs, float * dst, int64_t ds,
                           int N, int K, int M, void * stream);
/* debug hook: device buffer of (return value) int64 slots that CTA (0,0) of later b200_mul_mat_q launches fills with
 * clock64 stamps of its pipeline events (tools/mmq_debug.py trace); NULL switches it off. */
int b200_mmq_debug_trace(void * dev_buf);
/* MUL_MAT_ID: dst[:, e, t] = as[:, :, ids[e, t]] x b[:, e % ne11, t]; ids i32 read on the device (no host sync).
 * Replaces ggml_cuda_mul_mat_id (ggml-cuda/ggml-cuda.cu:2007-2148) and the MoE branch of mmq.cu:141-223. */
int b200_op_mul_mat_id(const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids, const b200_tensor * dst,
                       void * scratch, int64_t scratch_bytes, void * stream);
/* scratch the call above wants for these operands (0: the per-pair GEMV path needs none).  From 9 tokens on
 * (B200_MMQ_ID_MIN_T) the rows are sorted by expert on the device and multiplied by the tensor-core MMQ: */
int64_t b200_mul_mat_id_scratch_bytes(const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids);
/* the batched path itself.  W: n_expert matrices [N][K] of one quantised type, rows nb1 bytes apart, experts N * nb1 apart.
 * X: f32 activations of token t / slot u at X + t * x_sz + (u % x_ne1) * x_sy (floats).  ids: int32 [T][ids_s1], n_used
 * per token.  The result row of (t, u) goes to dst + t * d_sz + u * d_sy (floats).  Three launches: routing tables (counts,
 * offsets, (token, slot) pairs in expert order), gather + bf16 conversion, one MMQ grid over (token tiles, row tiles,
 * experts) whose CTAs leave when their expert has fewer tokens.  -1: shape not taken.  The reference sorts on the host after
 * a device->host copy of ids and a stream synchronisation (ggml-cuda.cu:2050-2075). */
int64_t b200_mul_mat_id_q_scratch_bytes(int K, int n_used, int T, int n_expert);
int b200_mul_mat_id_q(int type, const void * W, int64_t nb1, int n_expert, const float * X, int64_t x_sy, int64_t x_sz, int x_ne1,
                      const int32_t * ids, int64_t ids_s1, int n_used, int T, float * dst, int64_t d_sy, int64_t d_sz,
                      int N, int K, void * scratch, int64_t scratch_bytes, void * stream);

/* Fusions of adjacent graph nodes over rows of K floats (K % 32 == 0, rows 16-byte aligned; -1 = shape not taken, run the nodes
 * one by one).  Every intermediate tensor is written exactly as the separate operators write it.
 *   add_rms_norm_mul:  sum = a + b (skipped when b == NULL);  norm = rms_norm(sum or a, eps);  out = norm * w  (w: [K])
 *   unary_mul:         act = silu|gelu(g);  out = act * u
 * xq (optional): `out` once more in the form the following mat-mul reads, written by the same launch: xq_kind 0 = the GEMV's
 * pre-quantised activation layout (b200_quantized_acts_bytes; at most 8 rows), 1 = bf16 [M][K] for the MMQ (b200_acts_bf16_bytes).
 * Reference: the RMS_NORM+MUL(+ADD) and GLU fusions of ggml-cuda (ggml-cuda.cu ggml_cuda_can_fuse / norm.cu rms_norm_mul). */
int b200_op_add_rms_norm_mul(const b200_tensor * a, const b200_tensor * b, const b200_tensor * sum, const b200_tensor * norm,
                             const b200_tensor * w, const b200_tensor * out, float eps, void * xq, int xq_kind, void * stream);
int b200_op_unary_mul(int op, const b200_tensor * g, const b200_tensor * act, const b200_tensor * u, const b200_tensor * out,
                      void * xq, int xq_kind, void * stream);

/* Up to 4 rope / KV-store nodes of a decode step in one launch.
 *   kind 0: dst = rope(src)          kind 1: dst = rope(src) and dst2 = cpy(dst)  (the K row going into the cache)
 *   kind 2: dst2 = cpy(src)           (the V row going into the cache);  src, dst f32; dst2 f32 / f16 / bf16
 * dst2_slot (optional): device address of a pointer that replaces dst2->data when the kernel runs (see b200_op_cpy_indirect).
 * Outputs are bit-identical to b200_op_rope followed by b200_op_cpy. */
typedef struct b200_ew_desc {
    int kind;
    const b200_tensor * src, * dst, * dst2, * pos, * ffac;
    const b200_rope_params * rope;
    void * const * dst2_slot;
} b200_ew_desc;
int b200_op_rope_cpy_batch(const b200_ew_desc * tasks, int n, void * stream);
int b200_op_rope_cpy_batch_check(const b200_ew_desc * tasks, int n);      /* 0 when the batch would be accepted; launches nothing */

/* ---- the mat-vec chain of a decode step as ONE persistent launch (csrc/dchain.cu) -----------------------------------------
 * Between two attentions a batch-1 decode step of a Llama layer is
 *     wo mat-vec (+ residual) -> RMS_NORM * w -> gate / up mat-vecs -> act(gate) * up -> down mat-vec (+ residual)
 *     [ -> RMS_NORM * w -> up to two groups of mat-vecs on the same row: q / k / v of the next layer (rope + KV store), or lm_head ]
 * which the reference runs as 12-14 kernels (ggml-cuda/ggml-cuda.cu:2150-2400: mul_mat_vec_q + quantize_q8_1 per MUL_MAT, norm.cu,
 * binbcast.cu, unary.cu, rope.cu, cpy.cu).  Here the CTAs stay resident across the chain, the weight stream never stops at a node
 * boundary and the phases hand over through counters in L2.  Every tensor gets the bits the node-by-node entry points of this
 * header give it.  All weights Q4_K or Q6_K, K multiples of 256, norm rows <= 4096 floats, M = 1.
 *   gemv.epi: 0 dst[n] = row sum; 1 also sum[n] = dst[n] + resid[n] (ng = 1); 2 rope / KV-cache store as b200_mmvq_args.rope */
typedef struct b200_chain_gemv {
    int type, ng, K, epi;
    const void * W[3]; int N[3]; float * dst[3];
    const float * resid; float * sum;
    const struct b200_gemv_rope * rope;
} b200_chain_gemv;
typedef struct b200_chain_norm { const float * x; const float * w; float * y; float * out; float eps; int K; } b200_chain_norm;   /* y = rms_norm(x), out = y * w */
typedef struct b200_chain_args {
    b200_chain_gemv wo; const void * xq_attn;        /* xq_attn: the attention output as pre-quantised activations (b200_op_flash_attn_ext_xq) */
    b200_chain_norm norm1;                           /* x = wo.sum */
    b200_chain_gemv gate_up;                         /* ng = 2: dst[0] = gate, dst[1] = up (N[0] == N[1]) */
    int act_op; float * act; float * act_out;        /* act = unary(gate) (B200_UNARY_SILU / _GELU), act_out = act * up */
    void * xq_ffn;                                   /* scratch, b200_quantized_acts_bytes(N_ff, 1): act_out as the down projection reads it */
    b200_chain_gemv down;                            /* K = N_ff, resid = wo.sum */
    int tail;                                        /* 1: norm2 (x = down.sum) + t1 (+ t2 when t2.ng > 0, same activations) */
    b200_chain_norm norm2;
    b200_chain_gemv t1, t2;
    uint32_t * barriers;                             /* 128 zeroed 32-bit words of device memory; shared by every chain launch of a stream (they leave it zeroed) */
    uint32_t * error; unsigned timeout_ms;           /* bounded spins: *error = 2 on expiry (device-visible word, may be NULL), 0 = unbounded */
    int pdl;                                         /* as b200_mmvq_args.pdl */
    void * trace;                                    /* debug: 128 uint64 slots that receive %globaltimer stamps of the first and the last CTA (tools/chain_micro.py); NULL = off */
} b200_chain_args;
int b200_decode_chain_supported(const b200_chain_args * args);      /* host-side check only */
int b200_decode_chain(const b200_chain_args * args, void * stream);

/* number of kernels launched by this library so far in this process (bench.py gpu_launches) */
int b200_kernel_launch_count(void);
/* account for kernels launched through a replayed CUDA graph (the plugin captures decode steps) */
int b200_kernel_launch_add(int n);

/* ---- hand-over between devices through flags in peer memory (csrc/peer.cu) ----------------------------------
 * Replaces the cudaEventRecord / cudaStreamWaitEvent fork-join and the peer memcpys of the reference's split
 * MUL_MAT (ggml-cuda/ggml-cuda.cu:1605-1608, :1628-1630, :1701-1745).  All pointers are device pointers that the
 * launching device can reach (its own memory or peer-mapped memory); flags are 32-bit words, zero when idle.
 *   b200_peer_bcast : copy `bytes` (multiple of 16) from src to each dsts[p], then raise flags[p] (p < n_peers <=
 *                     B200_PEER_MAX); counter = one zeroed 32-bit word in the launching device's memory
 *   b200_peer_wait  : for every bit t of mask, wait until flags[t * stride] is raised and lower it (t < 32)
 *   b200_peer_signal: make this stream's earlier stores visible system-wide, then raise *remote_flag */
#define B200_PEER_MAX 15
int b200_peer_bcast(const void * src, void * const * dsts, uint32_t * const * flags, int n_peers, int64_t bytes,
                    unsigned * counter, void * stream);
int b200_peer_wait(uint32_t * flags, int stride, uint32_t mask, void * stream);
/* same with an explicit spin budget (0 = none) and an error word (device-visible, e.g. mapped host memory) that is set to 1
 * when the budget runs out; the kernel then carries on instead of trapping.  b200_peer_timeout_ms: env B200_SPLIT_TIMEOUT_MS,
 * default 30000. */
int b200_peer_wait_ex(uint32_t * flags, int stride, uint32_t mask, unsigned timeout_ms, uint32_t * error, void * stream);
unsigned b200_peer_timeout_ms(void);
int b200_peer_signal(uint32_t * remote_flag, void * stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_OPS_H */
