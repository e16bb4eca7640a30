"""GPU parity of MUL_MAT_ID (b200_op_mul_mat_id through the C ABI) against the CPU oracle.

Reference: ggml_cuda_mul_mat_id (ggml-cuda/ggml-cuda.cu:2007-2148) and the cases of tests/test-backend-ops.cpp:4297-4330
(n_mats 4/8, n_used 1..4, b broadcast or per-slot, n = 1..129 tokens).  A few tokens take one GEMV per (token, expert)
pair; from 9 tokens on the rows are sorted by expert on the device and go through the tensor-core MMQ.  Bar: NMSE <= 5e-4
against dequantise-then-f32 (the reference's MUL_MAT_ID bar); expert ids are integer work: a wrong route shows as NMSE ~ 1."""
import numpy as np
import pytest
import torch

import oracle as O

pytestmark = pytest.mark.gpu
NAMES = list(O.QUANT_TYPES)


def _expected(t, W, X, ids, N, K, n_expert):
    T, n_used = ids.shape
    rb = O.row_bytes(t, K)
    We = np.ascontiguousarray(W).reshape(n_expert, N * rb)
    want = np.zeros((T, n_used, N), np.float32)
    for e in range(n_expert):
        tt, uu = np.nonzero(ids == e)
        if tt.size == 0:
            continue
        rows = X[tt, uu % X.shape[1]]
        want[tt, uu] = O.mul_mat_f32(t, We[e], rows, N, K, tt.size)
    return want


def _run(t, W, X, ids, N, K, n_expert):
    from llamacog_b200 import ops
    flat = np.ascontiguousarray(W).reshape(-1)
    buf = torch.zeros(flat.size + 256, dtype=torch.uint8, device="cuda")
    buf[:flat.size] = torch.from_numpy(flat).cuda()
    got = ops.mul_mat_id(t, buf, torch.from_numpy(X).cuda(), torch.from_numpy(ids).cuda(), N, K, n_expert)
    torch.cuda.synchronize()
    return got.cpu().numpy()


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("n_expert,n_used,T,bcast", [(4, 2, 1, True), (8, 2, 5, False), (8, 2, 9, True), (8, 2, 12, False), (4, 1, 33, True),
                                                     (8, 4, 129, False), (8, 2, 300, True), (8, 2, 512, False), (3, 3, 40, True)])
def test_mul_mat_id_vs_oracle(name, n_expert, n_used, T, bcast):
    t = O.QUANT_TYPES[name]
    N, K = 160, 2048 if name == "q6_K" else 512
    rng = np.random.default_rng(n_expert * 100 + T)
    W = O.random_quant_bytes(t, n_expert * N, K, rng)
    X = rng.standard_normal((T, 1 if bcast else n_used, K)).astype(np.float32)
    ids = np.stack([rng.permutation(n_expert)[:n_used] for _ in range(T)]).astype(np.int32)     # distinct experts per token, as top-k gives
    got = _run(t, W, X, ids, N, K, n_expert)
    want = _expected(t, W, X, ids, N, K, n_expert)
    nmse = float(((got.astype(np.float64) - want) ** 2).sum() / (want.astype(np.float64) ** 2).sum())
    assert nmse < 5e-4, nmse
    # per row as well: one misrouted row would hide in the total
    err = ((got.astype(np.float64) - want) ** 2).sum(-1) / np.maximum((want.astype(np.float64) ** 2).sum(-1), 1e-30)
    assert err.max() < 5e-3, float(err.max())


def test_skewed_routing_and_empty_experts():
    """every token to the same two experts (the others get no rows), and repeated experts inside one token"""
    t = O.QUANT_TYPES["q4_K"]
    n_expert, n_used, T, N, K = 8, 2, 64, 256, 512
    rng = np.random.default_rng(5)
    W = O.random_quant_bytes(t, n_expert * N, K, rng)
    X = rng.standard_normal((T, 1, K)).astype(np.float32)
    for ids in (np.tile(np.array([[6, 1]], np.int32), (T, 1)), np.tile(np.array([[3, 3]], np.int32), (T, 1))):
        got = _run(t, W, X, ids, N, K, n_expert)
        want = _expected(t, W, X, ids, N, K, n_expert)
        nmse = float(((got.astype(np.float64) - want) ** 2).sum() / (want.astype(np.float64) ** 2).sum())
        assert nmse < 5e-4, nmse


def test_batched_equals_per_pair_rows(monkeypatch):
    """the two paths must agree row by row up to the activation rounding (bf16 vs q8_1): same routing, same weights"""
    t = O.QUANT_TYPES["q4_K"]
    n_expert, n_used, T, N, K = 8, 2, 48, 128, 1024
    rng = np.random.default_rng(9)
    W = O.random_quant_bytes(t, n_expert * N, K, rng)
    X = rng.standard_normal((T, 1, K)).astype(np.float32)
    ids = np.stack([rng.permutation(n_expert)[:n_used] for _ in range(T)]).astype(np.int32)
    a = _run(t, W, X, ids, N, K, n_expert)
    # (the threshold is read once per process: force the other path through its own entry point)
    from llamacog_b200 import ops
    b = np.stack([_run(t, W, X[i:i + 1], ids[i:i + 1], N, K, n_expert)[0] for i in range(0, T, 8)])
    a8 = a[::8]
    nmse = float(((a8.astype(np.float64) - b) ** 2).sum() / (b.astype(np.float64) ** 2).sum())
    assert nmse < 5e-4, nmse


@pytest.mark.parametrize("n_expert,n_used,T,n_embd,inplace", [(8, 2, 1, 4096, False), (8, 2, 1, 4096, True), (8, 2, 3, 512, False), (16, 4, 2, 300, False),
                                                              (32, 8, 1, 1000, True), (4, 3, 8, 64, False)])
def test_moe_combine_matches_single_operators(n_expert, n_used, T, n_embd, inplace):
    """b200_op_moe_combine (the tail of build_moe_ffn in one launch: GET_ROWS, SUM_ROWS, DIV, MUL, ADD x (n_used - 1)) must leave in
    every tensor the bits of the single operators run one after the other -- also with DIV in place on the weights, MUL in place on
    the experts and the ADDs on the running sum (what the graph allocator does)."""
    import ctypes as C
    from llamacog_b200 import ops
    L = ops.lib()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.manual_seed(n_expert * 100 + n_used * 10 + T)
    probs = torch.softmax(torch.randn(T, n_expert, device="cuda"), dim=-1).contiguous()
    order = torch.argsort(probs, dim=-1, descending=True).to(torch.int32).contiguous()        # [T, n_expert]: top-k = a view of its first columns
    ids = order[:, :n_used]
    experts = torch.randn(T, n_used, n_embd, device="cuda")
    B = lambda t: C.byref(ops.as_b200(t))
    # the single operators, as the plugin runs them node by node
    w_ref = torch.empty(T, n_used, 1, device="cuda")
    assert L.b200_op_get_rows(B(probs.view(T, n_expert, 1)), B(ids), B(w_ref), st) == 0
    s_ref = torch.empty(T, 1, device="cuda")
    assert L.b200_op_sum_rows(B(w_ref.view(T, n_used)), B(s_ref), st) == 0
    n_ref = torch.empty(T, n_used, device="cuda")
    assert L.b200_op_binary(C.c_int(3), B(w_ref.view(T, n_used)), B(s_ref), B(n_ref), st) == 0
    m_ref = torch.empty_like(experts)
    assert L.b200_op_binary(C.c_int(2), B(experts), B(n_ref.view(T, n_used, 1)), B(m_ref), st) == 0
    a_ref, acc = [], m_ref[:, 0, :]
    for k in range(1, n_used):
        o = torch.empty(T, n_embd, device="cuda")
        assert L.b200_op_binary(C.c_int(0), B(acc), B(m_ref[:, k, :]), B(o), st) == 0
        a_ref.append(o); acc = o
    torch.cuda.synchronize()
    # one launch
    ex = experts.clone()
    w = torch.empty(T, n_used, device="cuda"); s = torch.empty(T, device="cuda")
    wn = w if inplace else torch.empty(T, n_used, device="cuda")
    m = ex if inplace else torch.empty_like(ex)
    adds = []
    for k in range(1, n_used):
        adds.append(adds[-1] if (inplace and adds) else torch.empty(T, n_embd, device="cuda"))
    ap = (C.c_void_p * 7)(*([a.data_ptr() for a in adds] + [None] * (7 - len(adds))))
    assert L.b200_op_moe_combine(B(probs.view(T, n_expert, 1)), B(ids), B(ex), C.c_void_p(w.data_ptr()), C.c_void_p(s.data_ptr()), C.c_void_p(wn.data_ptr()),
                                 C.c_void_p(m.data_ptr()), ap, st) == 0
    torch.cuda.synchronize()
    assert torch.equal(s, s_ref.view(T)) and torch.equal(wn, n_ref)
    if not inplace:
        assert torch.equal(w, w_ref.view(T, n_used)) and torch.equal(m, m_ref)
        for a, r in zip(adds, a_ref):
            assert torch.equal(a, r)
    assert torch.equal(adds[-1], a_ref[-1])


@pytest.mark.parametrize("n_expert,T", [(8, 1), (8, 5), (16, 2), (60, 3), (128, 1), (1000, 2)])
def test_soft_max_argsort_matches_the_two_operators(n_expert, T):
    """The router's SOFT_MAX + ARGSORT (ggml_top_k) in one launch: probabilities and order bit-identical to b200_op_soft_max followed by
    b200_op_argsort, also with the probabilities written in place over the logits."""
    import ctypes as C
    from llamacog_b200 import ops
    L = ops.lib()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    torch.manual_seed(n_expert + T)
    logits = torch.randn(T, n_expert, device="cuda") * 3
    logits[0, : min(4, n_expert)] = 1.25                      # ties: the order of equal keys must match as well
    B = lambda t: C.byref(ops.as_b200(t))
    p_ref = torch.empty_like(logits); o_ref = torch.empty(T, n_expert, dtype=torch.int32, device="cuda")
    assert L.b200_op_soft_max(B(logits), None, B(p_ref), C.c_float(1.0), C.c_float(0.0), st) == 0
    assert L.b200_op_argsort(B(p_ref), B(o_ref), C.c_int(1), st) == 0
    for inplace in (False, True):
        x = logits.clone()
        p = x if inplace else torch.empty_like(x)
        o = torch.empty_like(o_ref)
        assert L.b200_op_soft_max_argsort(B(x), B(p), B(o), C.c_float(1.0), C.c_int(1), st) == 0
        torch.cuda.synchronize()
        assert torch.equal(p, p_ref) and torch.equal(o, o_ref)
