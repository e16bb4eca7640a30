"""GPU parity of the decode-step fusions (fused.cu) through the C ABI: every output of a fused launch must be bit-identical
to the outputs of the separate operators it replaces (b200_op_binary / rms_norm / unary / quantize_acts), which are themselves
checked against the reference by test-backend-ops and the oracle."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ops():
    from llamacog_b200 import ops
    return ops


def _st(ops):
    return ops._stream()


def _same_acts(q, q_ref):
    """the activation layout is 400 bytes per 256 values, of which the last 16 are padding"""
    M = q.shape[0]
    return torch.equal(q.view(M, -1, 400)[:, :, :384], q_ref.view(M, -1, 400)[:, :, :384])


@pytest.mark.parametrize("M", [1, 3, 8])
@pytest.mark.parametrize("K", [256, 2048, 4096, 8192, 14336])
@pytest.mark.parametrize("with_add", [False, True])
def test_add_rms_norm_mul_matches_separate_ops(M, K, with_add):
    ops = _ops(); L = ops.lib()
    torch.manual_seed(K + M)
    a = torch.randn(M, K, device="cuda"); b = torch.randn(M, K, device="cuda"); w = torch.randn(K, device="cuda")
    eps = 1e-5
    # separate operators
    s_ref = torch.empty_like(a); n_ref = torch.empty_like(a); o_ref = torch.empty_like(a)
    x = a
    if with_add:
        assert L.b200_op_binary(C.c_int(0), C.byref(ops.as_b200(a)), C.byref(ops.as_b200(b)), C.byref(ops.as_b200(s_ref)), _st(ops)) == 0
        x = s_ref
    assert L.b200_op_rms_norm(C.byref(ops.as_b200(x)), C.byref(ops.as_b200(n_ref)), C.c_float(eps), _st(ops)) == 0
    assert L.b200_op_binary(C.c_int(2), C.byref(ops.as_b200(n_ref)), C.byref(ops.as_b200(w)), C.byref(ops.as_b200(o_ref)), _st(ops)) == 0
    q_ref = ops.quantize_acts(o_ref)
    # fused
    s = torch.empty_like(a); n = torch.empty_like(a); o = torch.empty_like(a); q = torch.zeros_like(q_ref)
    rc = L.b200_op_add_rms_norm_mul(C.byref(ops.as_b200(a)), C.byref(ops.as_b200(b)) if with_add else None,
                                    C.byref(ops.as_b200(s)) if with_add else None, C.byref(ops.as_b200(n)), C.byref(ops.as_b200(w)),
                                    C.byref(ops.as_b200(o)), C.c_float(eps), C.c_void_p(q.data_ptr()), C.c_int(0), _st(ops))
    assert rc == 0
    torch.cuda.synchronize()
    if with_add:
        assert torch.equal(s, s_ref)
    assert torch.equal(n, n_ref) and torch.equal(o, o_ref)
    assert _same_acts(q, q_ref)
    # and against plain PyTorch fp32 (tolerance: fp32 summation order of the mean square)
    xx = (a + b) if with_add else a
    want = xx * torch.rsqrt((xx.double() ** 2).mean(dim=1, keepdim=True).float() + eps) * w
    assert torch.allclose(o, want, rtol=2e-5, atol=1e-6)


@pytest.mark.parametrize("M", [1, 5, 8])
@pytest.mark.parametrize("K", [512, 14336, 28672])
def test_silu_mul_matches_separate_ops(M, K):
    ops = _ops(); L = ops.lib()
    torch.manual_seed(K * 3 + M)
    g = torch.randn(M, K, device="cuda") * 3; u = torch.randn(M, K, device="cuda")
    a_ref = torch.empty_like(g); o_ref = torch.empty_like(g)
    assert L.b200_op_unary(C.c_int(10), C.byref(ops.as_b200(g)), C.byref(ops.as_b200(a_ref)), _st(ops)) == 0      # GGML_UNARY_OP_SILU
    assert L.b200_op_binary(C.c_int(2), C.byref(ops.as_b200(a_ref)), C.byref(ops.as_b200(u)), C.byref(ops.as_b200(o_ref)), _st(ops)) == 0
    q_ref = ops.quantize_acts(o_ref)
    a = torch.empty_like(g); o = torch.empty_like(g); q = torch.zeros_like(q_ref)
    assert L.b200_op_unary_mul(C.c_int(10), C.byref(ops.as_b200(g)), C.byref(ops.as_b200(a)), C.byref(ops.as_b200(u)), C.byref(ops.as_b200(o)),
                               C.c_void_p(q.data_ptr()), C.c_int(0), _st(ops)) == 0
    torch.cuda.synchronize()
    assert torch.equal(a, a_ref) and torch.equal(o, o_ref) and _same_acts(q, q_ref)
    assert torch.allclose(o, torch.nn.functional.silu(g) * u, rtol=1e-5, atol=1e-6)


def test_shapes_the_fusion_does_not_take_are_refused():
    ops = _ops(); L = ops.lib()
    a = torch.randn(9, 256, device="cuda"); w = torch.randn(256, device="cuda"); n = torch.empty_like(a); o = torch.empty_like(a)
    q = torch.zeros(9 * 400, dtype=torch.uint8, device="cuda")
    assert L.b200_op_add_rms_norm_mul(C.byref(ops.as_b200(a)), None, None, C.byref(ops.as_b200(n)), C.byref(ops.as_b200(w)),
                                      C.byref(ops.as_b200(o)), C.c_float(1e-5), C.c_void_p(q.data_ptr()), C.c_int(0), _st(ops)) == -1   # q8 layout: at most 8 rows
    a = torch.randn(4, 250, device="cuda"); w = torch.randn(250, device="cuda"); n = torch.empty_like(a); o = torch.empty_like(a)
    assert L.b200_op_add_rms_norm_mul(C.byref(ops.as_b200(a)), None, None, C.byref(ops.as_b200(n)), C.byref(ops.as_b200(w)),
                                      C.byref(ops.as_b200(o)), C.c_float(1e-5), None, C.c_int(0), _st(ops)) == -1       # K % 32 != 0


@pytest.mark.parametrize("M,K", [(9, 256), (100, 4096), (512, 4096), (300, 14336)])
def test_prefill_rows_with_bf16_activations(M, K):
    """many rows: the fused kernels also hand the MMQ its bf16 activations (bit-identical to b200_convert_acts_bf16 of the f32 output)"""
    ops = _ops(); L = ops.lib()
    torch.manual_seed(M + K)
    a = torch.randn(M, K, device="cuda"); b = torch.randn(M, K, device="cuda"); w = torch.randn(K, device="cuda")
    s = torch.empty_like(a); n = torch.empty_like(a); o = torch.empty_like(a); xb = torch.zeros(M, K, dtype=torch.bfloat16, device="cuda")
    assert L.b200_op_add_rms_norm_mul(C.byref(ops.as_b200(a)), C.byref(ops.as_b200(b)), C.byref(ops.as_b200(s)), C.byref(ops.as_b200(n)),
                                      C.byref(ops.as_b200(w)), C.byref(ops.as_b200(o)), C.c_float(1e-5), C.c_void_p(xb.data_ptr()), C.c_int(1), _st(ops)) == 0
    s_ref = torch.empty_like(a); n_ref = torch.empty_like(a); o_ref = torch.empty_like(a)
    assert L.b200_op_binary(C.c_int(0), C.byref(ops.as_b200(a)), C.byref(ops.as_b200(b)), C.byref(ops.as_b200(s_ref)), _st(ops)) == 0
    assert L.b200_op_rms_norm(C.byref(ops.as_b200(s_ref)), C.byref(ops.as_b200(n_ref)), C.c_float(1e-5), _st(ops)) == 0
    assert L.b200_op_binary(C.c_int(2), C.byref(ops.as_b200(n_ref)), C.byref(ops.as_b200(w)), C.byref(ops.as_b200(o_ref)), _st(ops)) == 0
    torch.cuda.synchronize()
    assert torch.equal(s, s_ref) and torch.equal(n, n_ref) and torch.equal(o, o_ref)
    assert torch.equal(xb, o_ref.to(torch.bfloat16))
    g = torch.randn(M, K, device="cuda") * 2; u = torch.randn(M, K, device="cuda")
    act = torch.empty_like(g); out = torch.empty_like(g); xb2 = torch.zeros(M, K, dtype=torch.bfloat16, device="cuda")
    assert L.b200_op_unary_mul(C.c_int(10), C.byref(ops.as_b200(g)), C.byref(ops.as_b200(act)), C.byref(ops.as_b200(u)), C.byref(ops.as_b200(out)),
                               C.c_void_p(xb2.data_ptr()), C.c_int(1), _st(ops)) == 0
    torch.cuda.synchronize()
    assert torch.allclose(out, torch.nn.functional.silu(g) * u, rtol=1e-5, atol=1e-6)
    assert torch.equal(xb2, out.to(torch.bfloat16))


def _gemv_args(ops, t, W, N, K, dst, **kw):
    return ops.MmvqArgs(type=t, W=W.data_ptr(), nb1=ops.row_bytes(t, K), X=None, xs=K, dst=dst.data_ptr(), ds=N, N=N, K=K, M=1,
                        ny=1, nz=1, pdl=-1, **kw)


@pytest.mark.parametrize("tname", ["q4_K", "q6_K", "q8_0", "q4_0", "q5_K", "q4_1", "q5_0", "q3_K", "iq4_xs"])
@pytest.mark.parametrize("N,K", [(4096, 4096), (1024, 2048), (300, 1024), (128256, 4096)])
def test_gemv_norm_prologue_matches_separate_launches(tname, N, K):
    """RMS_NORM + MUL + quantisation done by the batch-1 GEMV itself (b200_mmvq_args.norm_*): the normalised row, the
    product, the quantised column it leaves behind and the mat-vec result must equal, bit for bit, those of the stand-alone
    fused norm kernel followed by the GEMV on its activations."""
    import numpy as np
    import oracle as O
    ops = _ops(); L = ops.lib()
    t = O.ALL_QUANT_TYPES[tname]
    rng = np.random.default_rng(N + K)
    W = torch.from_numpy(O.random_quant_bytes(t, N, K, rng).reshape(-1)).cuda()
    torch.manual_seed(N + K)
    x = torch.randn(1, K, device="cuda") * 3; w = torch.randn(K, device="cuda")
    eps = 1e-5
    n_ref = torch.empty_like(x); o_ref = torch.empty_like(x); q_ref = torch.zeros((1, K // 256 * 400), dtype=torch.uint8, device="cuda")
    assert L.b200_op_add_rms_norm_mul(C.byref(ops.as_b200(x)), None, None, C.byref(ops.as_b200(n_ref)), C.byref(ops.as_b200(w)),
                                      C.byref(ops.as_b200(o_ref)), C.c_float(eps), C.c_void_p(q_ref.data_ptr()), C.c_int(0), _st(ops)) == 0
    d_ref = ops.mul_mat_vec_q(t, W, o_ref, N, K, Xq=q_ref)
    n = torch.empty_like(x); o = torch.empty_like(x); q = torch.zeros_like(q_ref); d = torch.empty(1, N, device="cuda")
    a = _gemv_args(ops, t, W, N, K, d, Xq=q.data_ptr(), norm_x=x.data_ptr(), norm_w=w.data_ptr(), norm_y=n.data_ptr(),
                   norm_out=o.data_ptr(), norm_eps=eps)
    if not L.b200_mmvq_fusable(C.byref(a), C.c_int(1), None, None):
        pytest.skip("shape not taken by the register-resident kernel")
    assert L.b200_mul_mat_vec_q_ex(C.byref(a), _st(ops)) == 0
    torch.cuda.synchronize()
    assert torch.equal(n, n_ref) and torch.equal(o, o_ref)
    assert _same_acts(q, q_ref)
    assert torch.equal(d, d_ref)


@pytest.mark.parametrize("tname", ["q4_K", "q6_K", "q8_0", "q4_0", "q5_K", "q4_1", "q5_0", "q3_K", "iq4_xs"])
@pytest.mark.parametrize("N,K", [(4096, 4096), (4096, 14336), (37, 2048), (5000, 4096)])
def test_gemv_residual_epilogue(tname, N, K):
    """ADD fused into the GEMV's store (b200_mmvq_args.fuse_resid / fuse_sum): dst as without it, fuse_sum = dst + resid exactly
    (one fp32 add per element, as b200_op_binary does); also with fuse_sum aliasing dst (the graph allocator's in-place ADD)."""
    import numpy as np
    import oracle as O
    ops = _ops(); L = ops.lib()
    t = O.ALL_QUANT_TYPES[tname]
    rng = np.random.default_rng(N * 3 + K)
    W = torch.from_numpy(O.random_quant_bytes(t, N, K, rng).reshape(-1)).cuda()
    torch.manual_seed(N)
    x = torch.randn(1, K, device="cuda"); r = torch.randn(1, N, device="cuda")
    xq = ops.quantize_acts(x)
    d_ref = ops.mul_mat_vec_q(t, W, x, N, K, Xq=xq)
    d = torch.empty(1, N, device="cuda"); s = torch.empty(1, N, device="cuda")
    a = _gemv_args(ops, t, W, N, K, d, Xq=xq.data_ptr(), fuse_resid=r.data_ptr(), fuse_sum=s.data_ptr())
    if not L.b200_mmvq_fusable(C.byref(a), C.c_int(1), None, None):
        pytest.skip("shape not taken by the register-resident kernel")
    assert L.b200_mul_mat_vec_q_ex(C.byref(a), _st(ops)) == 0
    torch.cuda.synchronize()
    assert torch.equal(d, d_ref)
    assert torch.equal(s, d_ref + r)
    d2 = torch.empty(1, N, device="cuda")
    a = _gemv_args(ops, t, W, N, K, d2, Xq=xq.data_ptr(), fuse_resid=r.data_ptr(), fuse_sum=d2.data_ptr())
    assert L.b200_mul_mat_vec_q_ex(C.byref(a), _st(ops)) == 0
    torch.cuda.synchronize()
    assert torch.equal(d2, d_ref + r)


@pytest.mark.parametrize("tname", ["q4_K", "q6_K", "q8_0", "q4_0", "q5_K", "q4_1", "q5_0", "q3_K", "iq4_xs"])
@pytest.mark.parametrize("N,K,inplace", [(14336, 4096, False), (14336, 4096, True), (28672, 4096, False), (2048, 2048, True), (96, 1024, False), (8192, 2048, False),
                                          (28672, 8192, True), (11008, 4096, False), (9472, 2048, True)])      # Llama-3-70B, Llama-2-7B (344 blocks: uneven CTAs), 296 blocks
def test_gemv_gate_up_silu_mul_epilogue(tname, N, K, inplace):
    """silu(gate) * up and its q8 activations computed by the gate / up group launch itself (b200_mmvq_args.sm_*, pair mode: every CTA
    owns the same 32-row blocks of both matrices).  gate, up, act, out and the quantised column must equal, bit for bit, the group
    launch followed by b200_op_unary_mul -- also with act / out written in place over gate / up (the graph allocator's choice), with the
    norm prologue in the same launch, and on a second launch over the same counters (they must come back to zero)."""
    import numpy as np
    import oracle as O
    ops = _ops(); L = ops.lib()
    t = O.ALL_QUANT_TYPES[tname]
    rng = np.random.default_rng(N + K)
    Wg = torch.from_numpy(O.random_quant_bytes(t, N, K, rng).reshape(-1)).cuda()
    Wu = torch.from_numpy(O.random_quant_bytes(t, N, K, rng).reshape(-1)).cuda()
    torch.manual_seed(N + K)
    x = torch.randn(1, K, device="cuda")
    xq = ops.quantize_acts(x)
    g_ref, u_ref = ops.mul_mat_vec_q_group(t, [Wg, Wu], [N, N], x, K, Xq=xq)
    a_ref = torch.empty_like(g_ref); o_ref = torch.empty_like(g_ref); q_ref = torch.zeros((1, (N + 255) // 256 * 400), dtype=torch.uint8, device="cuda")
    assert L.b200_op_unary_mul(C.c_int(10), C.byref(ops.as_b200(g_ref)), C.byref(ops.as_b200(a_ref)), C.byref(ops.as_b200(u_ref)), C.byref(ops.as_b200(o_ref)),
                               C.c_void_p(q_ref.data_ptr()), C.c_int(0), _st(ops)) == 0
    cnt = torch.zeros(2048, dtype=torch.int32, device="cuda")
    for rep in range(2):
        g = torch.empty(1, N, device="cuda"); u = torch.empty(1, N, device="cuda")
        act = g if inplace else torch.empty(1, N, device="cuda")
        out = u if inplace else torch.empty(1, N, device="cuda")
        q = torch.zeros_like(q_ref)
        a = _gemv_args(ops, t, Wg, N, K, g, Xq=xq.data_ptr(), sm_op=10, sm_act=act.data_ptr(), sm_out=out.data_ptr(), sm_xq=q.data_ptr(),
                       sm_counters=cnt.data_ptr())
        Wp = (C.c_void_p * 2)(Wg.data_ptr(), Wu.data_ptr()); Dp = (C.c_void_p * 2)(g.data_ptr(), u.data_ptr())
        Np = (C.c_int * 2)(N, N); dsp = (C.c_int64 * 2)(N, N)
        if not L.b200_mmvq_fusable(C.byref(a), C.c_int(2), Wp, Np):
            pytest.skip("shape not taken by the register-resident kernel")
        assert L.b200_mul_mat_vec_q_group(C.byref(a), C.c_int(2), Wp, Dp, Np, dsp, _st(ops)) == 0
        torch.cuda.synchronize()
        if not inplace:
            assert torch.equal(g, g_ref) and torch.equal(u, u_ref)
        assert torch.equal(act, a_ref) and torch.equal(out, o_ref)
        assert _same_acts(q, q_ref)
        assert int(cnt.abs().sum().item()) == 0


@pytest.mark.parametrize("tname", ["q4_K", "q6_K", "q8_0"])
@pytest.mark.parametrize("N,K,n_expert,n_used,inplace", [(14336, 4096, 8, 2, False), (14336, 4096, 8, 2, True), (512, 1024, 4, 3, False), (2048, 2048, 4, 1, True)])
def test_moe_gate_up_pair_launch(tname, N, K, n_expert, n_used, inplace):
    """The gate and up expert mat-vecs of one token, silu(gate) * up and its q8 activations in ONE launch (b200_mmvq_args.sm_up_*, pair
    mode over the chosen experts): every tensor equals, bit for bit, what the per-expert launches followed by b200_op_unary_mul leave."""
    import numpy as np
    import oracle as O
    ops = _ops(); L = ops.lib()
    t = O.ALL_QUANT_TYPES[tname]
    rng = np.random.default_rng(N + K + n_used)
    rb = ops.row_bytes(t, K)
    Wg = torch.from_numpy(O.random_quant_bytes(t, N * n_expert, K, rng).reshape(-1)).cuda()
    Wu = torch.from_numpy(O.random_quant_bytes(t, N * n_expert, K, rng).reshape(-1)).cuda()
    torch.manual_seed(N + K)
    x = torch.randn(1, K, device="cuda")
    xq = ops.quantize_acts(x)
    ids = torch.from_numpy(rng.choice(n_expert, n_used, replace=False).astype(np.int32)).cuda()
    g_ref = torch.empty(n_used, N, device="cuda"); u_ref = torch.empty(n_used, N, device="cuda")
    for y in range(n_used):
        e = int(ids[y].item())
        g_ref[y] = ops.mul_mat_vec_q(t, Wg[e * N * rb:(e + 1) * N * rb], x, N, K, Xq=xq)[0]
        u_ref[y] = ops.mul_mat_vec_q(t, Wu[e * N * rb:(e + 1) * N * rb], x, N, K, Xq=xq)[0]
    a_ref = torch.empty_like(g_ref); o_ref = torch.empty_like(g_ref)
    q_ref = torch.zeros((n_used, (N + 255) // 256 * 400), dtype=torch.uint8, device="cuda")
    assert L.b200_op_unary_mul(C.c_int(10), C.byref(ops.as_b200(g_ref)), C.byref(ops.as_b200(a_ref)), C.byref(ops.as_b200(u_ref)), C.byref(ops.as_b200(o_ref)),
                               C.c_void_p(q_ref.data_ptr()), C.c_int(0), _st(ops)) == 0
    for rep in range(2):
        g = torch.empty(n_used, N, device="cuda"); u = torch.empty(n_used, N, device="cuda")
        act = g if inplace else torch.empty_like(g)
        out = u if inplace else torch.empty_like(g)
        q = torch.zeros_like(q_ref)
        a = _gemv_args(ops, t, Wg, N, K, g, Xq=xq.data_ptr(), sm_op=10, sm_act=act.data_ptr(), sm_out=out.data_ptr(), sm_xq=q.data_ptr() if N % 256 == 0 else None,
                       sm_up_W=Wu.data_ptr(), sm_up_dst=u.data_ptr(), ids=ids.data_ptr(), ids_sz=n_used, w_sy=N * rb, d_sy=N, x_mod_y=1)
        a.ny = n_used
        if not L.b200_mmvq_fusable(C.byref(a), C.c_int(0), None, None):
            pytest.skip("shape not taken by the register-resident kernel")
        assert L.b200_mul_mat_vec_q_ex(C.byref(a), _st(ops)) == 0
        torch.cuda.synchronize()
        if not inplace:
            assert torch.equal(g, g_ref) and torch.equal(u, u_ref)
        assert torch.equal(act, a_ref) and torch.equal(out, o_ref)
        if N % 256 == 0:
            assert _same_acts(q, q_ref)


def test_gemv_fusion_refused_when_not_eligible():
    """M = 2 does not run on the register-resident kernel: the fusable predicate says no and a call with fused fields fails
    loudly (-4) instead of silently dropping the fused work."""
    import oracle as O
    import numpy as np
    ops = _ops(); L = ops.lib()
    t = O.QUANT_TYPES["q4_K"]
    N, K = 64, 4096
    W = torch.from_numpy(O.random_quant_bytes(t, N, K, np.random.default_rng(0)).reshape(-1)).cuda()
    x = torch.randn(2, K, device="cuda"); r = torch.randn(2, N, device="cuda"); d = torch.empty(2, N, device="cuda"); s = torch.empty(2, N, device="cuda")
    xq = ops.quantize_acts(x)
    a = _gemv_args(ops, t, W, N, K, d, Xq=xq.data_ptr(), fuse_resid=r.data_ptr(), fuse_sum=s.data_ptr())
    a.M = 2
    assert L.b200_mmvq_fusable(C.byref(a), C.c_int(1), None, None) == 0
    assert L.b200_mul_mat_vec_q_ex(C.byref(a), _st(ops)) == -4


class _RopeParams(C.Structure):
    _fields_ = [("n_dims", C.c_int), ("mode", C.c_int), ("n_ctx_orig", C.c_int), ("freq_base", C.c_float), ("freq_scale", C.c_float),
                ("ext_factor", C.c_float), ("attn_factor", C.c_float), ("beta_fast", C.c_float), ("beta_slow", C.c_float)]


class _GemvRope(C.Structure):
    _fields_ = [("mode", C.c_int * 3), ("out", C.c_void_p * 3), ("cell", C.c_void_p * 3), ("cell_slot", C.c_void_p * 3),
                ("cell_type", C.c_int * 3), ("head_dim", C.c_int * 3), ("pos", C.c_void_p), ("ffac", C.c_void_p), ("prm", _RopeParams)]


@pytest.mark.parametrize("tname", ["q4_K", "q6_K", "q8_0", "q4_0", "q5_K", "q4_1", "q5_0", "q3_K", "iq4_xs"])
@pytest.mark.parametrize("hd,n_dims,pos,ext,use_ff,slot", [(128, 128, 77, 0.0, False, False), (128, 128, 4095, 0.0, True, True),
                                                           (64, 32, 5, 0.0, False, False), (128, 128, 300, 1.0, False, True)])
def test_gemv_rope_epilogue_matches_rope_op(tname, hd, n_dims, pos, ext, use_ff, slot):
    """ROPE + KV-cache stores in the q/k/v mat-vec's epilogue (b200_gemv_rope): rope(q) as f32, rope(k) as f32 and as the f16
    cache row, v as the f16 cache row must equal, bit for bit, b200_op_rope (and a round-to-nearest f16 conversion) applied
    to the plain grouped mat-vec's outputs; also through a device pointer slot (CUDA-graph replay), with frequency factors,
    partial rotation (n_dims < head size) and YaRN mixing."""
    import numpy as np
    import oracle as O
    ops = _ops(); L = ops.lib()
    t = O.ALL_QUANT_TYPES[tname]
    K, nh, nkv = 2048, 12, 4
    Ns = [nh * hd, nkv * hd, nkv * hd]
    rng = np.random.default_rng(hd + pos)
    Ws = [torch.from_numpy(O.random_quant_bytes(t, n, K, rng).reshape(-1)).cuda() for n in Ns]
    torch.manual_seed(pos)
    X = torch.randn(1, K, device="cuda")
    xq = ops.quantize_acts(X)
    ref = ops.mul_mat_vec_q_group(t, Ws, Ns, X, K, Xq=xq)
    prm = _RopeParams(n_dims=n_dims, mode=0, n_ctx_orig=8192, freq_base=500000.0, freq_scale=1.0 if ext == 0.0 else 0.25, ext_factor=ext,
                      attn_factor=1.0, beta_fast=32.0, beta_slow=1.0)
    posd = torch.tensor([pos], dtype=torch.int32, device="cuda")
    ff = (torch.rand(n_dims // 2, device="cuda") + 0.5) if use_ff else None
    want = []
    for g in range(2):
        src = ref[g].view(1, Ns[g] // hd, hd).contiguous(); dst = torch.empty_like(src)
        assert L.b200_op_rope(C.byref(ops.as_b200(src)), C.byref(ops.as_b200(posd)), C.byref(ops.as_b200(ff)) if use_ff else None,
                              C.byref(ops.as_b200(dst)), C.byref(prm), _st(ops)) == 0
        want.append(dst.view(1, -1))
    # fused
    oq = torch.zeros(1, Ns[0], device="cuda"); ok_ = torch.zeros(1, Ns[1], device="cuda")
    kc = torch.zeros(Ns[1], dtype=torch.float16, device="cuda"); vc = torch.zeros(Ns[2], dtype=torch.float16, device="cuda")
    slots = torch.tensor([kc.data_ptr(), vc.data_ptr()], dtype=torch.int64, device="cuda")
    gr = _GemvRope()
    gr.mode[0], gr.mode[1], gr.mode[2] = 1, 2, 3
    gr.out[0], gr.out[1] = oq.data_ptr(), ok_.data_ptr()
    if slot:
        gr.cell_slot[1], gr.cell_slot[2] = slots.data_ptr(), slots.data_ptr() + 8
    else:
        gr.cell[1], gr.cell[2] = kc.data_ptr(), vc.data_ptr()
    gr.cell_type[1] = gr.cell_type[2] = 1
    gr.head_dim[0] = gr.head_dim[1] = hd
    gr.pos = posd.data_ptr(); gr.ffac = ff.data_ptr() if use_ff else None; gr.prm = prm
    raw = [torch.zeros(1, n, device="cuda") for n in Ns]
    a = ops.MmvqArgs(type=t, W=Ws[0].data_ptr(), nb1=ops.row_bytes(t, K), X=None, xs=K, dst=raw[0].data_ptr(), ds=Ns[0], N=Ns[0], K=K, M=1,
                     ny=1, nz=1, pdl=-1, Xq=xq.data_ptr(), rope=C.addressof(gr))
    Wp = (C.c_void_p * 3)(*[w.data_ptr() for w in Ws]); Dp = (C.c_void_p * 3)(*[d.data_ptr() for d in raw])
    Np = (C.c_int * 3)(*Ns); dsp = (C.c_int64 * 3)(*Ns)
    if not L.b200_mmvq_fusable(C.byref(a), C.c_int(3), Wp, Np):
        pytest.skip("shape not taken by the register-resident kernel")
    assert L.b200_mul_mat_vec_q_group(C.byref(a), C.c_int(3), Wp, Dp, Np, dsp, _st(ops)) == 0
    torch.cuda.synchronize()
    assert torch.equal(oq, want[0]) and torch.equal(ok_, want[1])
    assert torch.equal(kc, want[1].view(-1).half())
    assert torch.equal(raw[2], ref[2]) and torch.equal(vc, ref[2].view(-1).half())


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
@pytest.mark.parametrize("hd,nh,n_dims,mode,T", [(128, 32, 128, 0, 512), (128, 8, 128, 0, 512), (64, 12, 32, 0, 300), (128, 16, 128, 2, 256),
                                               (80, 5, 40, 2, 333)])
def test_rope_prefill_kernel_matches_decode_kernel(dtype, hd, nh, n_dims, mode, T):
    """b200_op_rope on a prefill batch takes the kernel that walks the heads per (token, pair) thread (cos/sin once per position
    and pair); it must equal, bit for bit, the per-element kernel that small batches take (here: the same rows in chunks of 8 tokens)."""
    ops = _ops(); L = ops.lib()
    torch.manual_seed(T + nh)
    src = torch.randn(T, nh, hd, device="cuda").to(dtype)
    pos = torch.arange(1000, 1000 + T, dtype=torch.int32, device="cuda")
    prm = _RopeParams(n_dims=n_dims, mode=mode, n_ctx_orig=8192, freq_base=500000.0, freq_scale=1.0, ext_factor=0.0, attn_factor=1.0,
                      beta_fast=32.0, beta_slow=1.0)
    got = torch.empty_like(src); want = torch.empty_like(src)
    assert L.b200_op_rope(C.byref(ops.as_b200(src)), C.byref(ops.as_b200(pos)), None, C.byref(ops.as_b200(got)), C.byref(prm), _st(ops)) == 0
    for t0 in range(0, T, 8):
        s = src[t0:t0 + 8]; d = want[t0:t0 + 8]; p = pos[t0:t0 + 8]
        assert L.b200_op_rope(C.byref(ops.as_b200(s)), C.byref(ops.as_b200(p)), None, C.byref(ops.as_b200(d)), C.byref(prm), _st(ops)) == 0
    torch.cuda.synchronize()
    assert torch.equal(got, want)
    if n_dims < hd:
        assert torch.equal(got[..., n_dims:], src[..., n_dims:])


@pytest.mark.parametrize("sd,dd", [(torch.float32, torch.float16), (torch.float32, torch.bfloat16), (torch.float16, torch.float32),
                                   (torch.float32, torch.float32)])
def test_cpy_vector_path_strided(sd, dd):
    """CPY of a prefill batch into a strided destination (the K / V store: rows of the batch into rows of a wider cache) and from a
    strided source, through the four-elements-per-thread kernel: exact round-to-nearest conversion, untouched bytes stay untouched."""
    ops = _ops(); L = ops.lib()
    torch.manual_seed(7)
    T, n = 512, 1024
    src_full = torch.randn(T, n + 64, device="cuda").to(sd)
    src = src_full[:, 32:32 + n]                      # strided source view, 16-byte aligned rows
    cache = torch.full((T + 8, 2 * n), 7.0, device="cuda").to(dd)
    dst = cache[4:4 + T, n:]                          # strided destination view
    assert L.b200_op_cpy(C.byref(ops.as_b200(src)), C.byref(ops.as_b200(dst)), _st(ops)) == 0
    # different logical shapes with the same element count (flat order): [T, n] -> [2T, n/2]
    flat = torch.zeros(2 * T, n // 2, device="cuda").to(dd)
    assert L.b200_op_cpy(C.byref(ops.as_b200(src)), C.byref(ops.as_b200(flat)), _st(ops)) == 0
    torch.cuda.synchronize()
    assert torch.equal(dst, src.to(dd))
    assert torch.equal(flat.view(T, n), src.to(dd))
    assert bool((cache[:4] == 7).all()) and bool((cache[4 + T:] == 7).all()) and bool((cache[:, :n] == 7).all())


class _EwDesc(C.Structure):
    _fields_ = [("kind", C.c_int), ("src", C.c_void_p), ("dst", C.c_void_p), ("dst2", C.c_void_p), ("pos", C.c_void_p), ("ffac", C.c_void_p),
                ("rope", C.c_void_p), ("dst2_slot", C.c_void_p)]


@pytest.mark.parametrize("T,cache_dtype,mode", [(512, torch.float16, 0), (300, torch.float16, 2), (512, torch.bfloat16, 0), (6, torch.float16, 0)])
def test_rope_store_batch_prefill_matches_separate_ops(T, cache_dtype, mode):
    """b200_op_rope_cpy_batch on a prefill batch (the wide kernel: 8 heads per thread, 4-wide V store) and on a decode batch: rope(q),
    rope(k) + K-cache rows, V-cache rows in one launch must equal b200_op_rope / b200_op_cpy run one after the other, bit for bit;
    the cache views are strided (rows of a larger cache)."""
    ops = _ops(); L = ops.lib()
    hd, nh, nkv, n_ctx = 128, 32, 8, 1024
    torch.manual_seed(T)
    q = torch.randn(T, nh, hd, device="cuda"); k = torch.randn(T, nkv, hd, device="cuda"); v = torch.randn(T, nkv * hd, device="cuda")
    pos = torch.arange(100, 100 + T, dtype=torch.int32, device="cuda")
    prm = _RopeParams(n_dims=hd, mode=mode, n_ctx_orig=8192, freq_base=500000.0, freq_scale=1.0, ext_factor=0.0, attn_factor=1.0, beta_fast=32.0, beta_slow=1.0)

    def run(batched):
        oq = torch.zeros_like(q); ok_ = torch.zeros_like(k)
        kc = torch.zeros(n_ctx, nkv * hd, device="cuda").to(cache_dtype); vc = torch.zeros(n_ctx, nkv * hd, device="cuda").to(cache_dtype)
        kview = kc[17:17 + T]; vview = vc[17:17 + T]
        tq, toq, tk, tok, tv, tpos = (ops.as_b200(x) for x in (q, oq, k, ok_, v, pos))
        tkc, tvc = ops.as_b200(kview), ops.as_b200(vview)
        if batched:
            d = (_EwDesc * 3)()
            d[0].kind = 0; d[0].src = C.addressof(tq); d[0].dst = C.addressof(toq); d[0].pos = C.addressof(tpos); d[0].rope = C.addressof(prm)
            d[1].kind = 1; d[1].src = C.addressof(tk); d[1].dst = C.addressof(tok); d[1].pos = C.addressof(tpos); d[1].rope = C.addressof(prm); d[1].dst2 = C.addressof(tkc)
            d[2].kind = 2; d[2].src = C.addressof(tv); d[2].dst2 = C.addressof(tvc)
            assert L.b200_op_rope_cpy_batch_check(d, 3) == 0
            assert L.b200_op_rope_cpy_batch(d, 3, _st(ops)) == 0
        else:
            assert L.b200_op_rope(C.byref(tq), C.byref(tpos), None, C.byref(toq), C.byref(prm), _st(ops)) == 0
            assert L.b200_op_rope(C.byref(tk), C.byref(tpos), None, C.byref(tok), C.byref(prm), _st(ops)) == 0
            tok2 = ops.as_b200(ok_.view(T, nkv * hd))
            assert L.b200_op_cpy(C.byref(tok2), C.byref(tkc), _st(ops)) == 0
            assert L.b200_op_cpy(C.byref(tv), C.byref(tvc), _st(ops)) == 0
        torch.cuda.synchronize()
        return oq, ok_, kc, vc

    a = run(True); b = run(False)
    for x, y, what in zip(a, b, ("rope(q)", "rope(k)", "K cache", "V cache")):
        assert torch.equal(x.view(torch.int16) if x.dtype != torch.float32 else x, y.view(torch.int16) if y.dtype != torch.float32 else y), what
    assert bool((a[2][:17] == 0).all()) and bool((a[2][17 + T:] == 0).all())
