"""ctypes bindings of libb200ops.so (include/b200_ops.h) over torch CUDA tensors.

Mirrors the reference operator semantics: `mul_mat(type, W, X)` is GGML_OP_MUL_MAT with a quantised
src0 (reference ggml-cuda/mmvq.cu:491-561), etc.  torch is used for device memory and streams only.
"""
import ctypes as C
import os

import torch

from . import build as _build

F32, F16, Q4_0, Q8_0, Q8_1, Q4_K, Q5_K, Q6_K, I32 = 0, 1, 2, 8, 9, 12, 13, 14, 26
Q4_1, Q5_0, Q5_1 = 3, 6, 7
Q2_K, Q3_K = 10, 11
IQ4_NL, IQ4_XS = 20, 23
BLOCK_BYTES = {IQ4_NL: 18, IQ4_XS: 136, Q2_K: 84, Q3_K: 110, Q4_1: 20, Q5_0: 22, Q5_1: 24, Q4_0: 18, Q8_0: 34, Q8_1: 36, Q4_K: 144, Q5_K: 176, Q6_K: 210, F32: 4, F16: 2}
BLOCK_ELEMS = {IQ4_NL: 32, IQ4_XS: 256, Q2_K: 256, Q3_K: 256, Q4_1: 32, Q5_0: 32, Q5_1: 32, Q4_0: 32, Q8_0: 32, Q8_1: 32, Q4_K: 256, Q5_K: 256, Q6_K: 256, F32: 1, F16: 1}


class MmvqArgs(C.Structure):
    _fields_ = [("type", C.c_int), ("W", C.c_void_p), ("nb1", C.c_int64), ("X", C.c_void_p), ("xs", C.c_int64),
                ("dst", C.c_void_p), ("ds", C.c_int64), ("N", C.c_int), ("K", C.c_int), ("M", C.c_int),
                ("ny", C.c_int), ("nz", C.c_int), ("w_sy", C.c_int64), ("w_sz", C.c_int64), ("w_div_y", C.c_int),
                ("w_div_z", C.c_int), ("x_sy", C.c_int64), ("x_sz", C.c_int64), ("x_mod_y", C.c_int),
                ("d_sy", C.c_int64), ("d_sz", C.c_int64), ("ids", C.c_void_p), ("ids_sz", C.c_int64),
                ("pdl", C.c_int), ("dbg_sumi", C.c_void_p), ("Xq", C.c_void_p), ("xq_sy", C.c_int64), ("xq_sz", C.c_int64),
                ("pf_n", C.c_int), ("pf_ptr", C.c_void_p * 3), ("pf_bytes", C.c_int64 * 3),
                ("fuse_resid", C.c_void_p), ("fuse_sum", C.c_void_p),
                ("norm_x", C.c_void_p), ("norm_w", C.c_void_p), ("norm_y", C.c_void_p), ("norm_out", C.c_void_p), ("norm_eps", C.c_float),
                ("rope", C.c_void_p),
                ("hs_wait", C.c_void_p), ("hs_clear", C.c_void_p), ("hs_signal", C.c_void_p), ("hs_join", C.c_void_p),
                ("hs_join_stride", C.c_int), ("hs_join_mask", C.c_uint32), ("hs_counter", C.c_void_p), ("hs_timeout_ms", C.c_uint),
                ("hs_error", C.c_void_p),
                ("sm_op", C.c_int), ("sm_act", C.c_void_p), ("sm_out", C.c_void_p), ("sm_xq", C.c_void_p), ("sm_counters", C.c_void_p),
                ("sm_up_W", C.c_void_p), ("sm_up_dst", C.c_void_p)]


_lib = None


def lib():
    """Load the CUDA library; fail loudly when it is missing (no CPU fallback exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(_build.OPS_LIB):
            raise RuntimeError("libb200ops.so is not built: run `python -m llamacog_b200.build` "
                               "(or __graft_entry__.build()); there is no CPU fallback")
        _lib = C.CDLL(_build.OPS_LIB)
        _lib.b200_mul_mat_vec_q_ex.argtypes = [C.POINTER(MmvqArgs), C.c_void_p]
        for name in EXPORTS:
            getattr(_lib, name).restype = C.c_int
        _lib.b200_quantized_acts_bytes.restype = C.c_int64
        _lib.b200_mul_mat_q_scratch_bytes.restype = C.c_int64
        _lib.b200_mul_mat_id_q_scratch_bytes.restype = C.c_int64
        _lib.b200_mul_mat_id_scratch_bytes.restype = C.c_int64
        _lib.b200_acts_bf16_bytes.restype = C.c_int64
        _lib.b200_flash_attn_scratch_bytes.restype = C.c_int64
    return _lib


# every symbol include/b200_ops.h declares (tests/test_capi_symbols.py keeps this in sync)
EXPORTS = ["b200_flash_attn_xq_supported", "b200_op_flash_attn_ext_xq", "b200_flash_attn_xb_supported", "b200_op_flash_attn_ext_xb", "b200_mul_mat_id_q", "b200_mul_mat_id_q_scratch_bytes", "b200_mul_mat_id_scratch_bytes", "b200_peer_bcast", "b200_peer_wait", "b200_peer_signal", "b200_mul_mat_vec_q_ex", "b200_mul_mat_vec_q", "b200_mul_mat_vec_q_group", "b200_mmvq_supported_type", "b200_quantize_q8_1",
           "b200_dequantize_rows", "b200_kernel_launch_count", "b200_kernel_launch_add", "b200_quantize_acts", "b200_quantized_acts_bytes",
           "b200_op_binary", "b200_op_unary", "b200_op_scale", "b200_op_rms_norm", "b200_op_rope", "b200_op_cpy", "b200_op_cpy_indirect", "b200_op_cpy_supported",
           "b200_op_get_rows", "b200_op_soft_max", "b200_op_argsort", "b200_op_sum_rows", "b200_op_moe_combine", "b200_op_soft_max_argsort", "b200_op_flash_attn_ext",
           "b200_flash_attn_scratch_bytes", "b200_flash_attn_supported", "b200_op_mul_mat", "b200_mul_mat_scratch_bytes",
           "b200_mul_mat_supported", "b200_op_mul_mat_id", "b200_mul_mat_q", "b200_mul_mat_q_scratch_bytes",
           "b200_mul_mat_q_supported", "b200_mmq_debug_trace", "b200_acts_bf16_bytes",
           "b200_convert_acts_bf16", "b200_mul_mat_q_bf16", "b200_mul_mat_q_bf16_group", "b200_op_add_rms_norm_mul", "b200_op_unary_mul", "b200_op_rope_cpy_batch", "b200_op_rope_cpy_batch_check",
           "b200_ktrace_enable", "b200_ktrace_dump", "b200_mmvq_fusable", "b200_mul_mat_q_noalloc", "b200_peer_wait_ex", "b200_peer_timeout_ms", "b200_decode_chain", "b200_decode_chain_supported"]


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _check(rc, what):
    if rc != 0:
        raise RuntimeError("%s failed: rc=%d" % (what, rc))


def row_bytes(t, k):
    assert k % BLOCK_ELEMS[t] == 0
    return k // BLOCK_ELEMS[t] * BLOCK_BYTES[t]


def quantize_acts(X):
    """X float32 [M, K] -> the GEMV's pre-quantised activation layout (uint8 [M, ceil(K/256)*400])."""
    assert X.is_cuda and X.dtype == torch.float32 and X.dim() == 2 and X.stride(1) == 1
    M, K = X.shape
    out = torch.empty((M, (K + 255) // 256 * 400), dtype=torch.uint8, device=X.device)
    _check(lib().b200_quantize_acts(C.c_void_p(X.data_ptr()), C.c_int64(X.stride(0)), C.c_void_p(out.data_ptr()),
                                    C.c_int(K), C.c_int(M), _stream()), "b200_quantize_acts")
    return out


def mul_mat_vec_q(t, W, X, N, K, nb1=None, dbg=False, pdl=-1, Xq=None):
    """W: uint8 CUDA tensor of N quantised rows; X: float32 [M, K].  Returns dst [M, N]
    (and the int32 block sums [M, N, K/32] when dbg).  Xq: activations pre-quantised by
    quantize_acts (the kernel then skips its fused prologue)."""
    assert W.is_cuda and X.is_cuda and X.dtype == torch.float32 and X.stride(-1) == 1
    M = X.shape[0]
    dst = torch.empty((M, N), dtype=torch.float32, device=X.device)
    sums = torch.zeros((M, N, K // 32), dtype=torch.int32, device=X.device) if dbg else None
    a = MmvqArgs(type=t, W=W.data_ptr(), nb1=nb1 or row_bytes(t, K), X=X.data_ptr(), xs=X.stride(0),
                 dst=dst.data_ptr(), ds=N, N=N, K=K, M=M, ny=1, nz=1, pdl=pdl, dbg_sumi=sums.data_ptr() if dbg else None,
                 Xq=Xq.data_ptr() if Xq is not None else None)
    _check(lib().b200_mul_mat_vec_q_ex(C.byref(a), _stream()), "b200_mul_mat_vec_q_ex")
    return (dst, sums) if dbg else dst


def mul_mat_vec_q_group(t, Ws, Ns, X, K, Xq=None, pdl=-1):
    """Up to three weight matrices of one type and K times the same activations in ONE launch (q/k/v, gate/up of a decode
    step; b200_mul_mat_vec_q_group).  Ws: uint8 CUDA tensors, Ns: their row counts; X float32 [M, K].  Returns a list of
    dst [M, N_g]."""
    assert X.is_cuda and X.dtype == torch.float32 and X.stride(-1) == 1 and 1 <= len(Ws) <= 3
    M, ng = X.shape[0], len(Ws)
    dsts = [torch.empty((M, n), dtype=torch.float32, device=X.device) for n in Ns]
    a = MmvqArgs(type=t, W=Ws[0].data_ptr(), nb1=row_bytes(t, K), X=X.data_ptr(), xs=X.stride(0), dst=dsts[0].data_ptr(), ds=Ns[0],
                 N=Ns[0], K=K, M=M, ny=1, nz=1, pdl=pdl, Xq=Xq.data_ptr() if Xq is not None else None)
    Wp = (C.c_void_p * ng)(*[w.data_ptr() for w in Ws])
    Dp = (C.c_void_p * ng)(*[d.data_ptr() for d in dsts])
    Np = (C.c_int * ng)(*Ns)
    dsp = (C.c_int64 * ng)(*Ns)
    _check(lib().b200_mul_mat_vec_q_group(C.byref(a), C.c_int(ng), Wp, Dp, Np, dsp, _stream()), "b200_mul_mat_vec_q_group")
    return dsts


def mul_mat_q(t, W, X, N, K, nb1=None):
    """Batched quantised mat-mul on the tensor cores (GGML_OP_MUL_MAT, quantised src0, M > 8 tokens).
    W: uint8 CUDA tensor of N quantised rows (16-byte aligned, a few bytes of slack after the last row);
    X: float32 [M, K].  Returns dst [M, N]."""
    assert W.is_cuda and X.is_cuda and X.dtype == torch.float32 and X.stride(-1) == 1
    M = X.shape[0]
    nb1 = nb1 or row_bytes(t, K)
    if not lib().b200_mul_mat_q_supported(C.c_int(t), C.c_void_p(W.data_ptr()), C.c_int64(nb1), C.c_int(K)):
        raise RuntimeError("b200_mul_mat_q: unsupported type/shape/alignment")
    need = lib().b200_mul_mat_q_scratch_bytes(C.c_int(K), C.c_int(M))
    scratch = torch.empty(need, dtype=torch.uint8, device=X.device)
    dst = torch.empty((M, N), dtype=torch.float32, device=X.device)
    _check(lib().b200_mul_mat_q(C.c_int(t), C.c_void_p(W.data_ptr()), C.c_int64(nb1), C.c_void_p(X.data_ptr()), C.c_int64(X.stride(0)),
                                C.c_void_p(dst.data_ptr()), C.c_int64(N), C.c_int(N), C.c_int(K), C.c_int(M),
                                C.c_void_p(scratch.data_ptr()), C.c_int64(need), _stream()), "b200_mul_mat_q")
    return dst


def mul_mat_q_noalloc(t, W, X, N, K, nb1=None, dst=None):
    """b200_mul_mat_q_noalloc: mul_mat_q with the activation scratch in the library's static device array (no allocation inside
    the call; pass `dst` to keep the binding allocation-free as well).  Returns (rc, dst): rc is the library's code."""
    assert W.is_cuda and X.is_cuda and X.dtype == torch.float32 and X.stride(-1) == 1
    M = X.shape[0]
    nb1 = nb1 or row_bytes(t, K)
    if dst is None:
        dst = torch.empty((M, N), dtype=torch.float32, device=X.device)
    rc = lib().b200_mul_mat_q_noalloc(C.c_int(t), C.c_void_p(W.data_ptr()), C.c_int64(nb1), C.c_void_p(X.data_ptr()), C.c_int64(X.stride(0)),
                                      C.c_void_p(dst.data_ptr()), C.c_int64(dst.stride(0)), C.c_int(N), C.c_int(K), C.c_int(M), _stream())
    return rc, dst


def mul_mat_q_group(t, Ws, Ns, X, K):
    """Several weight matrices of one type times the same activations in one tensor-core grid (b200_mul_mat_q_bf16_group)."""
    M = X.shape[0]
    nb1 = row_bytes(t, K)
    xb = torch.empty(lib().b200_acts_bf16_bytes(C.c_int(K), C.c_int(M)) + 128, dtype=torch.uint8, device=X.device)
    xp = (xb.data_ptr() + 127) & ~127
    _check(lib().b200_convert_acts_bf16(C.c_void_p(X.data_ptr()), C.c_int64(X.stride(0)), C.c_void_p(xp), C.c_int(K), C.c_int(M), _stream()),
           "b200_convert_acts_bf16")
    ng = len(Ws)
    outs = [torch.empty((M, n), dtype=torch.float32, device=X.device) for n in Ns]
    Wp = (C.c_void_p * ng)(*[w.data_ptr() for w in Ws]); Dp = (C.c_void_p * ng)(*[o.data_ptr() for o in outs])
    Np = (C.c_int * ng)(*Ns); dsp = (C.c_int64 * ng)(*Ns)
    _check(lib().b200_mul_mat_q_bf16_group(C.c_int(t), C.c_int(ng), Wp, C.c_int64(nb1), C.c_void_p(xp), Dp, dsp, Np, C.c_int(K), C.c_int(M),
                                           _stream()), "b200_mul_mat_q_bf16_group")
    return outs


def mul_mat_id(t, W, X, ids, N, K, n_expert):
    """GGML_OP_MUL_MAT_ID (quantised experts).  W: uint8 CUDA tensor holding n_expert matrices of N quantised rows each;
    X: float32 [T, x_ne1, K] with x_ne1 = 1 (shared by the token's experts) or n_used; ids: int32 [T, n_used].
    Returns dst [T, n_used, N].  From 9 tokens on the rows are sorted by expert and go through the tensor-core MMQ."""
    assert W.is_cuda and X.is_cuda and ids.is_cuda and X.dtype == torch.float32 and ids.dtype == torch.int32
    T, n_used = ids.shape
    nb1 = row_bytes(t, K)
    dst = torch.empty((T, n_used, N), dtype=torch.float32, device=X.device)
    a = B200Tensor(); a.data = W.data_ptr(); a.type = t
    a.ne[0], a.ne[1], a.ne[2], a.ne[3] = K, N, n_expert, 1
    a.nb[0], a.nb[1], a.nb[2], a.nb[3] = nb1 // (K // {2: 32, 8: 32}.get(t, 256)), nb1, nb1 * N, nb1 * N * n_expert
    b, i, d = as_b200(X), as_b200(ids), as_b200(dst)
    need = lib().b200_mul_mat_id_scratch_bytes(C.byref(a), C.byref(b), C.byref(i))
    scratch = torch.empty(max(need, 1), dtype=torch.uint8, device=X.device)
    _check(lib().b200_op_mul_mat_id(C.byref(a), C.byref(b), C.byref(i), C.byref(d), C.c_void_p(scratch.data_ptr() if need else None),
                                    C.c_int64(need), _stream()), "b200_op_mul_mat_id")
    return dst


def quantize_q8_1(X, k_pad=None):
    assert X.is_cuda and X.dtype == torch.float32 and X.dim() == 2 and X.stride(1) == 1
    rows, k = X.shape
    k_pad = k_pad or (k + 31) // 32 * 32
    y = torch.empty((rows, k_pad // 32 * 36), dtype=torch.uint8, device=X.device)
    _check(lib().b200_quantize_q8_1(C.c_void_p(X.data_ptr()), C.c_int64(X.stride(0)), C.c_void_p(y.data_ptr()),
                                    C.c_int64(k), C.c_int64(k_pad), C.c_int64(rows), _stream()), "b200_quantize_q8_1")
    return y


def dequantize_rows(t, W, nrows, k, out_dtype=torch.float32):
    assert W.is_cuda and W.dtype == torch.uint8
    dst = torch.empty((nrows, k), dtype=out_dtype, device=W.device)
    _check(lib().b200_dequantize_rows(C.c_int(t), C.c_void_p(W.data_ptr()), C.c_int64(row_bytes(t, k)),
                                      C.c_void_p(dst.data_ptr()), C.c_int(F16 if out_dtype == torch.float16 else F32),
                                      C.c_int64(k), C.c_int64(k), C.c_int64(nrows), _stream()), "b200_dequantize_rows")
    return dst


def kernel_launch_count():
    return lib().b200_kernel_launch_count()


class B200Tensor(C.Structure):
    """b200_tensor of include/b200_ops.h: data pointer, ggml type id, ne[4], nb[4] (bytes)."""
    _fields_ = [("data", C.c_void_p), ("type", C.c_int), ("ne", C.c_int64 * 4), ("nb", C.c_int64 * 4)]


_GGML_OF = {torch.float32: 0, torch.float16: 1, torch.int32: 26, torch.bfloat16: 30, torch.uint8: 24}       # uint8: raw blocks (I8 id as a placeholder; the caller sets the quantised type)


def as_b200(t):
    """A torch tensor (up to 4-D, torch order = ggml order reversed) as a b200_tensor view."""
    assert t.is_cuda and t.dim() <= 4
    shape = list(t.shape)[::-1] + [1] * (4 - t.dim())
    es = t.element_size()
    strides = [s * es for s in list(t.stride())[::-1]]
    while len(strides) < 4:
        strides.append(strides[-1] * shape[len(strides) - 1])
    r = B200Tensor()
    r.data = t.data_ptr(); r.type = _GGML_OF[t.dtype]
    for i in range(4):
        r.ne[i] = shape[i]; r.nb[i] = strides[i]
    return r


def flash_attn_ext(q, k, v, mask=None, scale=1.0, max_bias=0.0, logit_softcap=0.0):
    """GGML_OP_FLASH_ATTN_EXT.  q: f32 [ne3, n_head, N, D]; k, v: f16 [ne3, n_head_kv, n_kv, D]; mask: f16 [N_pad, n_kv_pad] or
    None.  Returns f32 [ne3, N, n_head, D] (the op's output layout)."""
    ne3, n_head, N, D = q.shape
    dst = torch.empty((ne3, N, n_head, D), dtype=torch.float32, device=q.device)
    tq, tk, tv, td = as_b200(q), as_b200(k), as_b200(v), as_b200(dst)
    for t, x in ((tk, k), (tv, v)):
        if x.dtype == torch.uint8:                   # q8_0 cache: uint8 [ne3, n_head_kv, n_kv, D / 32 * 34] holding block_q8_0 rows
            t.type = 8; t.ne[0] = D; t.nb[0] = 34
    tm = as_b200(mask) if mask is not None else None
    need = lib().b200_flash_attn_scratch_bytes(C.byref(tq), C.byref(tk))
    scratch = torch.empty(max(int(need), 16), dtype=torch.uint8, device=q.device)
    _check(lib().b200_op_flash_attn_ext(C.byref(tq), C.byref(tk), C.byref(tv), C.byref(tm) if tm is not None else None, C.byref(td),
                                        C.c_float(scale), C.c_float(max_bias), C.c_float(logit_softcap),
                                        C.c_void_p(scratch.data_ptr()), C.c_int64(scratch.numel()), _stream()), "b200_op_flash_attn_ext")
    return dst
