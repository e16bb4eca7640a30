#!/usr/bin/env python
"""Diagnostics for the tensor-core MMQ kernel: one-hot activations read the dequantised weights back
through the MMA (dst[m][n] = W'[n][k=m]), so a layout/swizzle/descriptor bug shows up as a pattern over
(n, k) instead of a single NMSE number.  Also times the kernel at model shapes.

usage: mmq_debug.py [check|time] [type ...]
"""
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import oracle as O  # noqa: E402  (diagnostic tool: the oracle is the checker)
from llamacog_b200 import ops  # noqa: E402


def dev_bytes(Wnp):
    flat = np.ascontiguousarray(Wnp).reshape(-1)
    buf = torch.zeros(flat.size + 256, dtype=torch.uint8, device="cuda")
    buf[:flat.size] = torch.from_numpy(flat).cuda()
    return buf


def onehot(name, N, K, tn=None):
    t = O.QUANT_TYPES[name]
    rng = np.random.default_rng(5)
    W = O.random_quant_bytes(t, N, K, rng)
    want = O.dequantize(t, W, N, K)                      # [N, K] f32
    X = torch.eye(K, dtype=torch.float32, device="cuda")  # M = K tokens, token m = e_m
    if tn:
        os.environ["B200_MMQ_TN"] = str(tn)
    got = ops.mul_mat_q(t, dev_bytes(W), X, N, K).cpu().numpy().T    # [N, K]
    torch.cuda.synchronize()
    err = np.abs(got - want)
    tol = np.abs(want) * 2.0 ** -6 + np.abs(want).max() * 2.0 ** -7
    bad = err > tol
    print("%-5s N=%d K=%d tn=%s: max|err|=%.3g  max|w|=%.3g  bad=%d/%d" % (name, N, K, tn, err.max(), np.abs(want).max(), bad.sum(), bad.size))
    if bad.any():
        rows = np.unique(np.nonzero(bad)[0]); cols = np.unique(np.nonzero(bad)[1])
        print("   bad rows (first 16):", rows[:16], " count", rows.size)
        print("   bad cols (first 32):", cols[:32], " count", cols.size)
        n, k = np.argwhere(bad)[0]
        print("   first bad: n=%d k=%d got=%g want=%g" % (n, k, got[n, k], want[n, k]))
        # is the value a known weight elsewhere in the row?  (tells a permutation from garbage)
        hits = np.nonzero(np.isclose(want[n], got[n, k], rtol=2e-2, atol=1e-6))[0]
        print("   got value matches want[n, k'] for k' in", hits[:8])
        print("   got[n, :16]  =", np.array2string(got[n, :16], precision=4))
        print("   want[n, :16] =", np.array2string(want[n, :16], precision=4))
    return not bad.any()


def nmse_case(name, N, K, M):
    t = O.QUANT_TYPES[name]
    rng = np.random.default_rng(N + K + M)
    W = O.random_quant_bytes(t, N, K, rng)
    Xn = rng.uniform(-1, 1, (M, K)).astype(np.float32)
    got = ops.mul_mat_q(t, dev_bytes(W), torch.from_numpy(Xn).cuda(), N, K).cpu().numpy()
    want = torch.from_numpy(O.dequantize(t, W, N, K)).cuda().double()
    ref = (torch.from_numpy(Xn).cuda().double() @ want.T).cpu().numpy()
    nmse = float(((got - ref) ** 2).sum() / (ref ** 2).sum())
    print("%-5s N=%d K=%d M=%d: nmse=%.3g" % (name, N, K, M, nmse))
    return nmse < 5e-4


def timing(name, N, K, M, iters=20):
    t = O.QUANT_TYPES[name]
    rb = ops.row_bytes(t, K)
    rng = np.random.default_rng(1)
    one = O.random_quant_bytes(t, 128, K, rng).reshape(-1)
    W = torch.from_numpy(np.tile(one, (N + 127) // 128)[: N * rb]).cuda()
    W = torch.cat([W, torch.zeros(256, dtype=torch.uint8, device="cuda")])
    X = torch.randn(M, K, device="cuda")
    for _ in range(3):
        ops.mul_mat_q(t, W, X, N, K)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        ops.mul_mat_q(t, W, X, N, K)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print("%-5s N=%5d K=%5d M=%4d tn=%s: %.3f ms  %.1f TFLOP/s" % (name, N, K, M, os.environ.get("B200_MMQ_TN", "auto"), ms, 2.0 * N * K * M / ms / 1e9))


def trace(name, N, K, M):
    """clock64 timeline of CTA (0,0): per 256-K block, when each role passed its waits (cycles since kernel start)."""
    import ctypes as C
    t = O.QUANT_TYPES[name]
    rb = ops.row_bytes(t, K)
    rng = np.random.default_rng(1)
    one = O.random_quant_bytes(t, 128, K, rng).reshape(-1)
    W = torch.from_numpy(np.tile(one, (N + 127) // 128)[: N * rb]).cuda()
    W = torch.cat([W, torch.zeros(256, dtype=torch.uint8, device="cuda")])
    X = torch.randn(M, K, device="cuda")
    for _ in range(2):
        ops.mul_mat_q(t, W, X, N, K)
    n = ops.lib().b200_mmq_debug_trace(None)
    buf = torch.zeros(n, dtype=torch.int64, device="cuda")
    ops.lib().b200_mmq_debug_trace(C.c_void_p(buf.data_ptr()))
    ops.mul_mat_q(t, W, X, N, K)
    torch.cuda.synchronize()
    ops.lib().b200_mmq_debug_trace(None)
    T = buf.cpu().numpy().reshape(24, -1)
    nkb = K // 256
    t0 = T[0, 0]
    rel = lambda v: int(v - t0) if v else -1
    print("trace %s N=%d K=%d M=%d tn=%s  (cycles since first W issue)" % (name, N, K, M, os.environ.get("B200_MMQ_TN", "auto")))
    print("kb | Wissue | g0: start rawfull emit0 aempty0 st0 emit1 aempty1 st1 | g1: start rawfull emit0 aempty0 st0 emit1 aempty1 st1")
    for kb in range(min(nkb, 16)):
        g = lambda h: " ".join("%6d" % rel(T[e + 8 * h, kb]) for e in (8, 9, 10, 11, 12, 13, 14, 15))
        print("%2d | %6d | %s | %s" % (kb, rel(T[0, kb]), g(0), g(1)))
    print("step | Xissue | mma: loop afull bfull issued")
    for i in range(min(4 * nkb, 48)):
        print("%3d | %6d | %6d %6d %6d %6d" % (i, rel(T[1, i]), rel(T[2, i]), rel(T[3, i]), rel(T[4, i]), rel(T[5, i])))
    print("tmem_full seen %d, kernel end %d" % (rel(T[6, 0]), rel(T[7, 0])))


def main():
    mode = sys.argv[1] if len(sys.argv) > 1 else "check"
    names = sys.argv[2:] or ["q4_K", "q6_K", "q5_K", "q4_0", "q8_0"]
    ok = True
    if mode == "check":
        for name in names:
            K = 2048 if name == "q6_K" else 256
            for tn in (32, 128, 256):
                ok &= onehot(name, 128, K, tn)
            os.environ.pop("B200_MMQ_TN", None)
            ok &= onehot(name, 200, 2048, None)
            ok &= nmse_case(name, 300, 4096, 77)
            ok &= nmse_case(name, 1024, 2048, 512)
        print("ALL OK" if ok else "FAILURES")
        sys.exit(0 if ok else 1)
    elif mode == "trace":
        for name in names:
            for tn in ("256", "128"):
                os.environ["B200_MMQ_TN"] = tn
                trace(name, int(os.environ.get("TRACE_N", "14336")), 4096, 512)
    else:
        for name in names:
            for (N, K, M) in [(1024, 4096, 512), (4096, 4096, 512), (14336, 4096, 512), (4096, 14336, 512), (4096, 4096, 128), (14336, 4096, 2048)]:
                for tn in ("128", "256"):
                    os.environ["B200_MMQ_TN"] = tn
                    timing(name, N, K, M)


if __name__ == "__main__":
    main()
