#!/usr/bin/env python
"""Tiny driver for ncu: a few direct launches of the decode GEMV on chosen shapes (no CUDA graph)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from llamacog_b200 import ops  # noqa: E402
from tools.mmvq_micro import rand_weights  # noqa: E402

SHAPES = {"down4k": (ops.Q4_K, 4096, 14336), "up4k": (ops.Q4_K, 14336, 4096), "q4k": (ops.Q4_K, 4096, 4096),
          "head6k": (ops.Q6_K, 128256, 4096), "down6k": (ops.Q6_K, 4096, 14336), "kv4k": (ops.Q4_K, 1024, 4096)}


def pair_launch(dev, N=14336, K=4096):
    """The gate / up launch of a decode step as the plugin issues it: pair mode with silu(gate) * up and its q8 activations
    (b200_mmvq_args.sm_*), three times over distinct weights."""
    import ctypes as C
    t = ops.Q4_K
    L = ops.lib()
    X = torch.rand((1, K), device=dev) * 2 - 1
    xq = ops.quantize_acts(X)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for _ in range(3):
        Wg = rand_weights(t, N * ops.row_bytes(t, K), dev); Wu = rand_weights(t, N * ops.row_bytes(t, K), dev)
        g = torch.empty(1, N, device=dev); u = torch.empty(1, N, device=dev); act = torch.empty(1, N, device=dev); out = torch.empty(1, N, device=dev)
        q = torch.zeros(N // 256 * 400, dtype=torch.uint8, device=dev)
        a = ops.MmvqArgs(type=t, W=Wg.data_ptr(), nb1=ops.row_bytes(t, K), X=None, xs=K, dst=g.data_ptr(), ds=N, N=N, K=K, M=1, ny=1, nz=1, pdl=-1,
                         Xq=xq.data_ptr(), sm_op=10, sm_act=act.data_ptr(), sm_out=out.data_ptr(), sm_xq=q.data_ptr())
        Wp = (C.c_void_p * 2)(Wg.data_ptr(), Wu.data_ptr()); Dp = (C.c_void_p * 2)(g.data_ptr(), u.data_ptr())
        Np = (C.c_int * 2)(N, N); dsp = (C.c_int64 * 2)(N, N)
        assert L.b200_mul_mat_vec_q_group(C.byref(a), C.c_int(2), Wp, Dp, Np, dsp, st) == 0
    torch.cuda.synchronize()


def main():
    dev = torch.device("cuda:0")
    names = sys.argv[1:] or list(SHAPES)
    if "pair4k" in names:
        pair_launch(dev)
        names = [n for n in names if n != "pair4k"]
    for n in names:
        t, N, K = SHAPES[n]
        pool = [rand_weights(t, N * ops.row_bytes(t, K), dev) for _ in range(3)]
        X = torch.rand((1, K), device=dev) * 2 - 1
        xq_once = ops.quantize_acts(X) if os.environ.get("MICRO_XQ", "0") == "2" else None
        for i in range(3):
            if xq_once is not None:      # activations quantised by their producer: the batch-1 register-resident kernel
                ops.mul_mat_vec_q(t, pool[i], X, N, K, Xq=xq_once)
            elif os.environ.get("MICRO_XQ", "0") == "1":
                ops.mul_mat_vec_q(t, pool[i], X, N, K, Xq=ops.quantize_acts(X))
            else:
                ops.mul_mat_vec_q(t, pool[i], X, N, K)
        torch.cuda.synchronize()
    print("done")


if __name__ == "__main__":
    main()
