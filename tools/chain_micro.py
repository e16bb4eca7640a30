#!/usr/bin/env python
"""Chain vs node-by-node timing of the decode mat-vec chain (no attention in between): L layers of distinct weights (> L2) replayed
from a CUDA graph.  usage: python tools/chain_micro.py [layers]"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from llamacog_b200 import ops  # noqa: E402
from tools.mmvq_micro import rand_weights  # noqa: E402
import test_gpu_chain as T  # noqa: E402

Q4_K, Q6_K = 12, 14
n_embd, n_ff, n_head, n_kv, hd = 4096, 14336, 32, 8, 128
L = int(sys.argv[1]) if len(sys.argv) > 1 else 8
dev = torch.device("cuda")
lib = ops.lib()


def W(t, N, K):
    return rand_weights(t, N * ops.row_bytes(t, K) + 256, dev)


layers = []
for il in range(L):
    more = il % 3 == 2
    layers.append(dict(wo=W(Q4_K, n_embd, n_embd), wg=W(Q4_K, n_ff, n_embd), wu=W(Q4_K, n_ff, n_embd), wd=W(Q6_K if more else Q4_K, n_embd, n_ff), td=Q6_K if more else Q4_K,
                       wq=W(Q4_K, n_embd, n_embd), wk=W(Q4_K, n_kv * hd, n_embd), wv=W(Q6_K if more else Q4_K, n_kv * hd, n_embd), tv=Q6_K if more else Q4_K))
attn = torch.randn(1, n_embd, device=dev); xq_attn = ops.quantize_acts(attn)
inp = torch.randn(1, n_embd, device=dev); w1 = torch.randn(n_embd, device=dev); w2 = torch.randn(n_embd, device=dev)
f = lambda n: torch.empty(1, n, device=dev)
d_o, s1, n1, o1, g, u, act, y, d_d, s2, n2, o2 = f(n_embd), f(n_embd), f(n_embd), f(n_embd), f(n_ff), f(n_ff), f(n_ff), f(n_ff), f(n_embd), f(n_embd), f(n_embd), f(n_embd)
sq, sk, v, rq, rk = f(n_embd), f(n_kv * hd), f(n_kv * hd), f(n_embd), f(n_kv * hd)
kc = torch.zeros(n_kv * hd, dtype=torch.float16, device=dev); vc = torch.zeros(n_kv * hd, dtype=torch.float16, device=dev)
xq1 = torch.zeros(n_embd // 256 * 400, dtype=torch.uint8, device=dev); xq2 = torch.zeros(n_ff // 256 * 400, dtype=torch.uint8, device=dev)
xq3 = torch.zeros_like(xq1)
posd = torch.tensor([345], dtype=torch.int32, device=dev)
prm = T.RopeParams(n_dims=hd, mode=0, n_ctx_orig=8192, freq_base=500000.0, freq_scale=1.0, ext_factor=0.0, attn_factor=1.0, beta_fast=32.0, beta_slow=1.0)
bars = torch.zeros(128, dtype=torch.int32, device=dev); err = torch.zeros(8, dtype=torch.int32, device=dev)
keep = []


def st():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def rope_desc(same):
    gr = T.GemvRope()
    gr.mode[0], gr.mode[1], gr.mode[2] = 1, 2, 3
    gr.out[0], gr.out[1] = rq.data_ptr(), rk.data_ptr()
    gr.cell[1], gr.cell[2] = kc.data_ptr(), vc.data_ptr()
    gr.cell_type[1] = gr.cell_type[2] = 1
    gr.head_dim[0] = gr.head_dim[1] = hd
    gr.pos = posd.data_ptr(); gr.prm = prm
    grv = T.GemvRope(); grv.mode[0] = 3; grv.cell[0] = vc.data_ptr(); grv.cell_type[0] = 1; grv.pos = posd.data_ptr(); grv.prm = prm
    keep.extend([gr, grv])
    return gr, grv


def run_nodes():
    for l in layers:
        a = ops.MmvqArgs(type=Q4_K, W=l["wo"].data_ptr(), nb1=ops.row_bytes(Q4_K, n_embd), X=None, xs=n_embd, dst=d_o.data_ptr(), ds=n_embd, N=n_embd, K=n_embd, M=1, ny=1, nz=1,
                         pdl=-1, Xq=xq_attn.data_ptr(), fuse_resid=inp.data_ptr(), fuse_sum=s1.data_ptr())
        assert lib.b200_mul_mat_vec_q_ex(C.byref(a), st()) == 0
        # gate/up with the norm in the prologue
        a = ops.MmvqArgs(type=Q4_K, W=l["wg"].data_ptr(), nb1=ops.row_bytes(Q4_K, n_embd), X=None, xs=n_embd, dst=g.data_ptr(), ds=n_ff, N=n_ff, K=n_embd, M=1, ny=1, nz=1, pdl=-1,
                         Xq=xq1.data_ptr(), norm_x=s1.data_ptr(), norm_w=w1.data_ptr(), norm_y=n1.data_ptr(), norm_out=o1.data_ptr(), norm_eps=1e-5)
        Wp = (C.c_void_p * 2)(l["wg"].data_ptr(), l["wu"].data_ptr()); Dp = (C.c_void_p * 2)(g.data_ptr(), u.data_ptr())
        Np = (C.c_int * 2)(n_ff, n_ff); dsp = (C.c_int64 * 2)(n_ff, n_ff)
        assert lib.b200_mul_mat_vec_q_group(C.byref(a), 2, Wp, Dp, Np, dsp, st()) == 0
        assert lib.b200_op_unary_mul(C.c_int(10), C.byref(ops.as_b200(g)), C.byref(ops.as_b200(act)), C.byref(ops.as_b200(u)), C.byref(ops.as_b200(y)),
                                     C.c_void_p(xq2.data_ptr()), C.c_int(0), st()) == 0
        a = ops.MmvqArgs(type=l["td"], W=l["wd"].data_ptr(), nb1=ops.row_bytes(l["td"], n_ff), X=None, xs=n_ff, dst=d_d.data_ptr(), ds=n_embd, N=n_embd, K=n_ff, M=1, ny=1, nz=1,
                         pdl=-1, Xq=xq2.data_ptr(), fuse_resid=s1.data_ptr(), fuse_sum=s2.data_ptr())
        assert lib.b200_mul_mat_vec_q_ex(C.byref(a), st()) == 0
        same = l["tv"] == Q4_K
        gr, grv = rope_desc(same)
        a = ops.MmvqArgs(type=Q4_K, W=l["wq"].data_ptr(), nb1=ops.row_bytes(Q4_K, n_embd), X=None, xs=n_embd, dst=sq.data_ptr(), ds=n_embd, N=n_embd, K=n_embd, M=1, ny=1, nz=1, pdl=-1,
                         Xq=xq3.data_ptr(), norm_x=s2.data_ptr(), norm_w=w2.data_ptr(), norm_y=n2.data_ptr(), norm_out=o2.data_ptr(), norm_eps=1e-5, rope=C.addressof(gr))
        ng = 3 if same else 2
        Wp = (C.c_void_p * 3)(l["wq"].data_ptr(), l["wk"].data_ptr(), l["wv"].data_ptr()); Dp = (C.c_void_p * 3)(sq.data_ptr(), sk.data_ptr(), v.data_ptr())
        Np = (C.c_int * 3)(n_embd, n_kv * hd, n_kv * hd); dsp = (C.c_int64 * 3)(n_embd, n_kv * hd, n_kv * hd)
        assert lib.b200_mul_mat_vec_q_group(C.byref(a), ng, Wp, Dp, Np, dsp, st()) == 0
        if not same:
            a = ops.MmvqArgs(type=Q6_K, W=l["wv"].data_ptr(), nb1=ops.row_bytes(Q6_K, n_embd), X=None, xs=n_embd, dst=v.data_ptr(), ds=n_kv * hd, N=n_kv * hd, K=n_embd, M=1, ny=1, nz=1,
                             pdl=-1, Xq=xq3.data_ptr(), rope=C.addressof(grv))
            assert lib.b200_mul_mat_vec_q_ex(C.byref(a), st()) == 0


def run_chain():
    for l in layers:
        a = T.ChainArgs()
        a.wo.type, a.wo.ng, a.wo.K, a.wo.epi = Q4_K, 1, n_embd, 1
        a.wo.W[0], a.wo.N[0], a.wo.dst[0] = l["wo"].data_ptr(), n_embd, d_o.data_ptr()
        a.wo.resid, a.wo.sum = inp.data_ptr(), s1.data_ptr()
        a.xq_attn = xq_attn.data_ptr()
        a.norm1 = T.ChainNorm(x=s1.data_ptr(), w=w1.data_ptr(), y=n1.data_ptr(), out=o1.data_ptr(), eps=1e-5, K=n_embd)
        a.gate_up.type, a.gate_up.ng, a.gate_up.K, a.gate_up.epi = Q4_K, 2, n_embd, 0
        a.gate_up.W[0], a.gate_up.W[1] = l["wg"].data_ptr(), l["wu"].data_ptr()
        a.gate_up.N[0] = a.gate_up.N[1] = n_ff
        a.gate_up.dst[0], a.gate_up.dst[1] = g.data_ptr(), u.data_ptr()
        a.act_op, a.act, a.act_out, a.xq_ffn = 10, act.data_ptr(), y.data_ptr(), xq2.data_ptr()
        a.down.type, a.down.ng, a.down.K, a.down.epi = l["td"], 1, n_ff, 1
        a.down.W[0], a.down.N[0], a.down.dst[0] = l["wd"].data_ptr(), n_embd, d_d.data_ptr()
        a.down.resid, a.down.sum = s1.data_ptr(), s2.data_ptr()
        a.tail = 1
        a.norm2 = T.ChainNorm(x=s2.data_ptr(), w=w2.data_ptr(), y=n2.data_ptr(), out=o2.data_ptr(), eps=1e-5, K=n_embd)
        same = l["tv"] == Q4_K
        gr, grv = rope_desc(same)
        a.t1.type, a.t1.ng, a.t1.K, a.t1.epi = Q4_K, 3 if same else 2, n_embd, 2
        a.t1.W[0], a.t1.W[1], a.t1.W[2] = l["wq"].data_ptr(), l["wk"].data_ptr(), l["wv"].data_ptr()
        a.t1.N[0], a.t1.N[1], a.t1.N[2] = n_embd, n_kv * hd, n_kv * hd
        a.t1.dst[0], a.t1.dst[1], a.t1.dst[2] = sq.data_ptr(), sk.data_ptr(), v.data_ptr()
        a.t1.rope = C.addressof(gr)
        if not same:
            a.t2.type, a.t2.ng, a.t2.K, a.t2.epi = Q6_K, 1, n_embd, 2
            a.t2.W[0], a.t2.N[0], a.t2.dst[0] = l["wv"].data_ptr(), n_kv * hd, v.data_ptr()
            a.t2.rope = C.addressof(grv)
        a.barriers = bars.data_ptr(); a.error = err.data_ptr(); a.timeout_ms = 2000; a.pdl = -1
        assert lib.b200_decode_chain(C.byref(a), st()) == 0


def timeit(fn, name):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn(); torch.cuda.synchronize()
        gph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gph, stream=s):
            fn()
        for _ in range(5):
            gph.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(20):
            gph.replay()
        e1.record(s)
        torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / 20 / L
    nbytes = sum(sum(int(t.numel()) - 256 for k, t in l.items() if k.startswith("w")) for l in layers) / L
    print("%-14s %7.2f us per layer   %.0f GB/s (%.1f MB per layer)   err=%d" % (name, us, nbytes / us / 1e3, nbytes / 1e6, int(err[0])))


timeit(run_nodes, "node by node")
timeit(run_chain, "chain")

# phase timeline of ONE chain launch (first and last CTA), after a warm launch
trace = torch.zeros(128, dtype=torch.int64, device=dev)
TRACE = [None]
_old = run_chain


def one_chain_traced():
    global layers
    keep_layers = layers
    layers = layers[:2]
    import types
    orig = T.ChainArgs
    # patch: set trace on the second launch only
    count = [0]
    real = lib.b200_decode_chain

    class Wrap:
        def __call__(self, aref, stream):
            count[0] += 1
            if count[0] == 2:
                aref._obj.trace = trace.data_ptr()
            return real(aref, stream)
    lib_b = Wrap()
    globals()["lib"] = type("L", (), {"b200_decode_chain": lib_b, "__getattr__": lambda self, n: getattr(ops.lib(), n)})()
    run_chain()
    torch.cuda.synchronize()
    globals()["lib"] = ops.lib()
    layers = keep_layers


one_chain_traced()
tr = trace.cpu().numpy().astype(np.int64)
names = {0: "start", 28: "silu: gate/up complete (barrier 1 passed)", 29: "silu: my blocks done"}
for ph, nm in enumerate(["wo", "gate/up", "down", "q/k", "v"]):
    names[1 + 4 * ph] = nm + ": wait passed"; names[2 + 4 * ph] = nm + ": activations in registers"
    names[3 + 4 * ph] = nm + ": first stage landed"; names[4 + 4 * ph] = nm + ": last stage multiplied"
    names[32 + 2 * ph] = nm + ": [w0] first stage epilogue"; names[33 + 2 * ph] = nm + ": [w0] last epilogue + arrive"
for nn in range(2):
    names[44 + 4 * nn] = "norm %d: row in shared memory" % nn; names[45 + 4 * nn] = "norm %d: sums of squares done (barrier)" % nn
    names[52 + 4 * nn] = "norm %d: partials summed" % nn; names[53 + 4 * nn] = "norm %d: scale known" % nn
    names[54 + 4 * nn] = "norm %d: first pass of the quantiser done" % nn
    names[46 + 4 * nn] = "norm %d: thread 0 quantised its blocks" % nn; names[47 + 4 * nn] = "norm %d: all threads done (barrier)" % nn
for base, who in ((0, "CTA 0"), (64, "last CTA")):
    t0 = tr[base + 0]
    ev = sorted((int(tr[base + i]), i) for i in range(64) if tr[base + i] != 0)
    print("---", who)
    for t, i in ev:
        print("  %8.2f us  %s" % ((t - t0) / 1e3, names.get(i, str(i))))
