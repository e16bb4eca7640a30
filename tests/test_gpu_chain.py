"""b200_decode_chain (csrc/dchain.cu): the mat-vec chain of a decode step between two attentions as ONE persistent launch, through
the C ABI, against the same chain run node by node through the stand-alone entry points of the library (each of which is held
to the oracle by test_gpu_mmvq.py / test_gpu_fused.py): wo + residual, RMS_NORM * w, gate / up, SILU * up, down + residual,
RMS_NORM * w, q / k / v with rope and KV-cache stores (or lm_head).

Bar: EVERY tensor the chain writes equals the node-by-node result bit for bit (integer block sums are order independent, the fp32
orders are those of the stand-alone kernels), twice in a row on the same barrier counters (they must return to zero), and the
error word stays clear.  Reference counterpart: the per-node loop of ggml_cuda_compute_forward (ggml-cuda/ggml-cuda.cu:2150-2400)."""
import ctypes as C

import numpy as np
import pytest
import torch

import oracle as O

pytestmark = pytest.mark.gpu


class RopeParams(C.Structure):
    _fields_ = [("n_dims", C.c_int), ("mode", C.c_int), ("n_ctx_orig", C.c_int), ("freq_base", C.c_float), ("freq_scale", C.c_float),
                ("ext_factor", C.c_float), ("attn_factor", C.c_float), ("beta_fast", C.c_float), ("beta_slow", C.c_float)]


class GemvRope(C.Structure):
    _fields_ = [("mode", C.c_int * 3), ("out", C.c_void_p * 3), ("cell", C.c_void_p * 3), ("cell_slot", C.c_void_p * 3),
                ("cell_type", C.c_int * 3), ("head_dim", C.c_int * 3), ("pos", C.c_void_p), ("ffac", C.c_void_p), ("prm", RopeParams)]


class ChainGemv(C.Structure):
    _fields_ = [("type", C.c_int), ("ng", C.c_int), ("K", C.c_int), ("epi", C.c_int), ("W", C.c_void_p * 3), ("N", C.c_int * 3),
                ("dst", C.c_void_p * 3), ("resid", C.c_void_p), ("sum", C.c_void_p), ("rope", C.c_void_p)]


class ChainNorm(C.Structure):
    _fields_ = [("x", C.c_void_p), ("w", C.c_void_p), ("y", C.c_void_p), ("out", C.c_void_p), ("eps", C.c_float), ("K", C.c_int)]


class ChainArgs(C.Structure):
    _fields_ = [("wo", ChainGemv), ("xq_attn", C.c_void_p), ("norm1", ChainNorm), ("gate_up", ChainGemv), ("act_op", C.c_int),
                ("act", C.c_void_p), ("act_out", C.c_void_p), ("xq_ffn", C.c_void_p), ("down", ChainGemv), ("tail", C.c_int),
                ("norm2", ChainNorm), ("t1", ChainGemv), ("t2", ChainGemv), ("barriers", C.c_void_p), ("error", C.c_void_p),
                ("timeout_ms", C.c_uint), ("pdl", C.c_int), ("trace", C.c_void_p)]


def _ops():
    from llamacog_b200 import ops
    return ops


def _st(ops):
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _w(t, N, K, rng):
    flat = O.random_quant_bytes(t, N, K, rng).reshape(-1)
    buf = torch.zeros(flat.size + 256, dtype=torch.uint8, device="cuda")
    buf[:flat.size] = torch.from_numpy(flat).cuda()
    return buf


def _gemv(ops, t, W, N, K, xq, dst, **kw):
    a = ops.MmvqArgs(type=t, W=W.data_ptr(), nb1=ops.row_bytes(t, K), X=None, xs=K, dst=dst.data_ptr(), ds=N, N=N, K=K, M=1, ny=1, nz=1, pdl=-1,
                     Xq=xq.data_ptr(), **kw)
    assert ops.lib().b200_mul_mat_vec_q_ex(C.byref(a), _st(ops)) == 0


def _norm(ops, x, w, eps):
    L = ops.lib()
    K = x.shape[1]
    n = torch.empty_like(x); o = torch.empty_like(x); q = torch.zeros((1, K // 256 * 400), dtype=torch.uint8, device="cuda")
    assert L.b200_op_add_rms_norm_mul(C.byref(ops.as_b200(x)), None, None, C.byref(ops.as_b200(n)), C.byref(ops.as_b200(w)), C.byref(ops.as_b200(o)),
                                      C.c_float(eps), C.c_void_p(q.data_ptr()), C.c_int(0), _st(ops)) == 0
    return n, o, q


def _same_acts(a, b):
    """quantised activation columns: compare the 48 meaningful bytes of every 32-block (items carry 16 bytes of padding)"""
    a = a.view(-1, 400)[:, :384]; b = b.view(-1, 400)[:, :384]
    return torch.equal(a, b)


CASES = [
    # n_embd, n_ff, n_head, n_kv, hd, type of (wo, gate/up, down, q/k, v), tail
    (4096, 14336, 32, 8, 128, ("q4_K", "q4_K", "q6_K", "q4_K", "q6_K"), "qkv"),
    (4096, 14336, 32, 8, 128, ("q4_K", "q4_K", "q4_K", "q4_K", "q4_K"), "qkv"),
    (4096, 11264, 32, 4, 128, ("q6_K", "q4_K", "q4_K", "q6_K", "q6_K"), "qkv"),
    (4096, 14336, 32, 8, 128, ("q4_K", "q4_K", "q6_K", "q6_K", None), "head"),
    (4096, 14336, 64, 8, 64, ("q4_K", "q6_K", "q6_K", "q4_K", "q4_K"), "none"),
]


@pytest.mark.parametrize("n_embd,n_ff,n_head,n_kv,hd,types,tail", CASES)
def test_chain_equals_node_by_node(n_embd, n_ff, n_head, n_kv, hd, types, tail):
    ops = _ops(); L = ops.lib()
    T = [O.QUANT_TYPES[x] if x else None for x in types]
    rng = np.random.default_rng(n_embd + n_ff + len(tail))
    torch.manual_seed(n_ff)
    eps = 1e-5
    Kq = n_head * hd
    Wo, Wg, Wu, Wd = _w(T[0], n_embd, Kq, rng), _w(T[1], n_ff, n_embd, rng), _w(T[1], n_ff, n_embd, rng), _w(T[2], n_embd, n_ff, rng)
    attn = torch.randn(1, Kq, device="cuda"); inp = torch.randn(1, n_embd, device="cuda")
    w1 = torch.randn(n_embd, device="cuda"); w2 = torch.randn(n_embd, device="cuda")
    xq_attn = ops.quantize_acts(attn)
    n_vocab = 32000
    if tail == "qkv":
        Nq, Nk = n_head * hd, n_kv * hd
        Wq, Wk, Wv = _w(T[3], Nq, n_embd, rng), _w(T[3], Nk, n_embd, rng), _w(T[4], Nk, n_embd, rng)
        prm = RopeParams(n_dims=hd, mode=0, n_ctx_orig=8192, freq_base=500000.0, freq_scale=1.0, ext_factor=0.0, attn_factor=1.0, beta_fast=32.0, beta_slow=1.0)
        posd = torch.tensor([345], dtype=torch.int32, device="cuda")
    elif tail == "head":
        Wh = _w(T[3], n_vocab, n_embd, rng)

    # ---- node by node ----
    r = {}
    r["d_o"] = torch.empty(1, n_embd, device="cuda"); r["s1"] = torch.empty(1, n_embd, device="cuda")
    _gemv(ops, T[0], Wo, n_embd, Kq, xq_attn, r["d_o"], fuse_resid=inp.data_ptr(), fuse_sum=r["s1"].data_ptr())
    r["n1"], r["o1"], xq1 = _norm(ops, r["s1"], w1, eps)
    r["g"], r["u"] = ops.mul_mat_vec_q_group(T[1], [Wg, Wu], [n_ff, n_ff], r["o1"], n_embd, Xq=xq1)
    r["act"] = torch.empty_like(r["g"]); r["y"] = torch.empty_like(r["g"]); xq2 = torch.zeros((1, n_ff // 256 * 400), dtype=torch.uint8, device="cuda")
    assert L.b200_op_unary_mul(C.c_int(10), C.byref(ops.as_b200(r["g"])), C.byref(ops.as_b200(r["act"])), C.byref(ops.as_b200(r["u"])), C.byref(ops.as_b200(r["y"])),
                               C.c_void_p(xq2.data_ptr()), C.c_int(0), _st(ops)) == 0
    r["d_d"] = torch.empty(1, n_embd, device="cuda"); r["s2"] = torch.empty(1, n_embd, device="cuda")
    _gemv(ops, T[2], Wd, n_embd, n_ff, xq2, r["d_d"], fuse_resid=r["s1"].data_ptr(), fuse_sum=r["s2"].data_ptr())
    if tail != "none":
        r["n2"], r["o2"], xq3 = _norm(ops, r["s2"], w2, eps)
    if tail == "qkv":
        rawq = torch.empty(1, Nq, device="cuda"); rawk = torch.empty(1, Nk, device="cuda"); r["v"] = torch.empty(1, Nk, device="cuda")
        _gemv(ops, T[3], Wq, Nq, n_embd, xq3, rawq); _gemv(ops, T[3], Wk, Nk, n_embd, xq3, rawk); _gemv(ops, T[4], Wv, Nk, n_embd, xq3, r["v"])
        for name, raw, nh in (("rq", rawq, n_head), ("rk", rawk, n_kv)):
            src = raw.view(1, nh, hd).contiguous(); dst = torch.empty_like(src)
            assert L.b200_op_rope(C.byref(ops.as_b200(src)), C.byref(ops.as_b200(posd)), None, C.byref(ops.as_b200(dst)), C.byref(prm), _st(ops)) == 0
            r[name] = dst.view(1, -1)
        r["kc"] = r["rk"].view(-1).half(); r["vc"] = r["v"].view(-1).half()
    elif tail == "head":
        r["logits"] = torch.empty(1, n_vocab, device="cuda")
        _gemv(ops, T[3], Wh, n_vocab, n_embd, xq3, r["logits"])
    torch.cuda.synchronize()

    # ---- the chain ----
    bars = torch.zeros(128, dtype=torch.int32, device="cuda"); err = torch.zeros(8, dtype=torch.int32, device="cuda")
    for rep in range(2):
        c = {k: torch.full_like(v, float("nan")) if v.dtype == torch.float32 else torch.zeros_like(v) for k, v in r.items()}
        xq_ffn = torch.zeros((1, n_ff // 256 * 400), dtype=torch.uint8, device="cuda")
        a = ChainArgs()
        a.wo.type, a.wo.ng, a.wo.K, a.wo.epi = T[0], 1, Kq, 1
        a.wo.W[0], a.wo.N[0], a.wo.dst[0] = Wo.data_ptr(), n_embd, c["d_o"].data_ptr()
        a.wo.resid, a.wo.sum = inp.data_ptr(), c["s1"].data_ptr()
        a.xq_attn = xq_attn.data_ptr()
        a.norm1 = ChainNorm(x=c["s1"].data_ptr(), w=w1.data_ptr(), y=c["n1"].data_ptr(), out=c["o1"].data_ptr(), eps=eps, K=n_embd)
        a.gate_up.type, a.gate_up.ng, a.gate_up.K, a.gate_up.epi = T[1], 2, n_embd, 0
        a.gate_up.W[0], a.gate_up.W[1] = Wg.data_ptr(), Wu.data_ptr()
        a.gate_up.N[0] = a.gate_up.N[1] = n_ff
        a.gate_up.dst[0], a.gate_up.dst[1] = c["g"].data_ptr(), c["u"].data_ptr()
        a.act_op, a.act, a.act_out, a.xq_ffn = 10, c["act"].data_ptr(), c["y"].data_ptr(), xq_ffn.data_ptr()
        a.down.type, a.down.ng, a.down.K, a.down.epi = T[2], 1, n_ff, 1
        a.down.W[0], a.down.N[0], a.down.dst[0] = Wd.data_ptr(), n_embd, c["d_d"].data_ptr()
        a.down.resid, a.down.sum = c["s1"].data_ptr(), c["s2"].data_ptr()
        a.tail = 0 if tail == "none" else 1
        gr = GemvRope(); grv = GemvRope()
        if tail != "none":
            a.norm2 = ChainNorm(x=c["s2"].data_ptr(), w=w2.data_ptr(), y=c["n2"].data_ptr(), out=c["o2"].data_ptr(), eps=eps, K=n_embd)
        if tail == "qkv":
            same = T[3] == T[4]
            sq = torch.empty(1, Nq, device="cuda"); sk = torch.empty(1, Nk, device="cuda")      # raw q / k: scratch (the rope outputs are what the graph keeps)
            gr.mode[0], gr.mode[1], gr.mode[2] = 1, 2, 3
            gr.out[0], gr.out[1] = c["rq"].data_ptr(), c["rk"].data_ptr()
            gr.cell[1], gr.cell[2] = c["kc"].data_ptr(), c["vc"].data_ptr()
            gr.cell_type[1] = gr.cell_type[2] = 1
            gr.head_dim[0] = gr.head_dim[1] = hd
            gr.pos = posd.data_ptr(); gr.prm = prm
            a.t1.type, a.t1.ng, a.t1.K, a.t1.epi = T[3], 3 if same else 2, n_embd, 2
            a.t1.W[0], a.t1.W[1], a.t1.W[2] = Wq.data_ptr(), Wk.data_ptr(), Wv.data_ptr()
            a.t1.N[0], a.t1.N[1], a.t1.N[2] = Nq, Nk, Nk
            a.t1.dst[0], a.t1.dst[1], a.t1.dst[2] = sq.data_ptr(), sk.data_ptr(), c["v"].data_ptr()
            a.t1.rope = C.addressof(gr)
            if not same:
                grv.mode[0] = 3; grv.cell[0] = c["vc"].data_ptr(); grv.cell_type[0] = 1; grv.pos = posd.data_ptr(); grv.prm = prm
                a.t2.type, a.t2.ng, a.t2.K, a.t2.epi = T[4], 1, n_embd, 2
                a.t2.W[0], a.t2.N[0], a.t2.dst[0] = Wv.data_ptr(), Nk, c["v"].data_ptr()
                a.t2.rope = C.addressof(grv)
        elif tail == "head":
            a.t1.type, a.t1.ng, a.t1.K, a.t1.epi = T[3], 1, n_embd, 0
            a.t1.W[0], a.t1.N[0], a.t1.dst[0] = Wh.data_ptr(), n_vocab, c["logits"].data_ptr()
        a.barriers = bars.data_ptr(); a.error = err.data_ptr(); a.timeout_ms = 2000; a.pdl = -1
        assert L.b200_decode_chain_supported(C.byref(a)) == 1
        assert L.b200_decode_chain(C.byref(a), _st(ops)) == 0
        torch.cuda.synchronize()
        assert int(err[0]) == 0, "a grid barrier of the chain timed out"
        for k in r:
            assert torch.equal(c[k], r[k]), (rep, k, int((c[k] != r[k]).sum()), c[k].numel())
        assert _same_acts(xq_ffn, xq2)
    # the counters close their ring: after a launch only the last one still holds its count, and the next launch clears it
    left = bars.view(4, 32)[:, 0].tolist()
    assert left[:3] == [0, 0, 0], left
