#!/usr/bin/env python
"""Micro-benchmark of the decode GEMV on the real Llama-3-8B shapes.
Each timed launch reads a DIFFERENT weight matrix from a pool larger than L2 (126 MB), so every
byte comes from HBM.  Prints one line per shape: us/launch, algorithmic GB/s, fraction of the
measured copy peak (MEASURED_PEAKS.json)."""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from llamacog_b200 import ops  # noqa: E402

SHAPES = [("attn_q/o 4096x4096 q4_K", ops.Q4_K, 4096, 4096), ("attn_k/v 1024x4096 q4_K", ops.Q4_K, 1024, 4096),
          ("ffn_up/gate 14336x4096 q4_K", ops.Q4_K, 14336, 4096), ("ffn_down 4096x14336 q4_K", ops.Q4_K, 4096, 14336),
          ("ffn_down 4096x14336 q6_K", ops.Q6_K, 4096, 14336), ("attn_v 1024x4096 q6_K", ops.Q6_K, 1024, 4096),
          ("lm_head 128256x4096 q6_K", ops.Q6_K, 128256, 4096),
          ("ffn_up 14336x4096 q4_0", ops.Q4_0, 14336, 4096), ("ffn_up 14336x4096 q8_0", ops.Q8_0, 14336, 4096),
          ("ffn_up 14336x4096 q5_K", ops.Q5_K, 14336, 4096), ("tbo-perf 4096x14336 q4_0", ops.Q4_0, 4096, 14336)]


def peak():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured"
    except Exception:
        return 6650.0, "fallback"


def rand_weights(t, nbytes, dev):
    """random payload with sane fp16 scales, generated on device (bench input only)."""
    bb = ops.BLOCK_BYTES[t]
    nb = nbytes // bb
    w = torch.randint(0, 256, (nb, bb), dtype=torch.uint8, device=dev)
    h = (torch.rand(nb, device=dev) * 0.01 + 0.002).to(torch.float16).view(torch.uint8).reshape(nb, 2)
    if t in (ops.Q4_0, ops.Q8_0):
        w[:, 0:2] = h
    elif t in (ops.Q4_K, ops.Q5_K):
        w[:, 0:2] = h
        w[:, 2:4] = h
    else:
        w[:, 208:210] = h
    return w.reshape(-1)


def main():
    dev = torch.device("cuda:0")
    pk, how = peak()
    Ms = [int(a) for a in sys.argv[1:]] or [1]
    for M in Ms:
        for name, t, N, K in SHAPES:
            if os.environ.get('MICRO_SHAPES') and not any(k in name for k in os.environ['MICRO_SHAPES'].split(',')):
                continue
            wbytes = N * ops.row_bytes(t, K)
            copies = max(2, int(400e6 // wbytes) + 1)
            pool = [rand_weights(t, wbytes, dev) for _ in range(min(copies, 48))]
            X = torch.rand((M, K), device=dev) * 2 - 1
            prequant = os.environ.get("MICRO_XQ", "0") == "1"
            xq_once = ops.quantize_acts(X) if os.environ.get("MICRO_XQ", "0") == "2" else None

            def run(Wm):
                if xq_once is not None:   # activations quantised once by their producer (the decode step's normal case)
                    return ops.mul_mat_vec_q(t, Wm, X, N, K, Xq=xq_once)
                if prequant:   # separate quantise kernel + GEMV on pre-quantised activations
                    return ops.mul_mat_vec_q(t, Wm, X, N, K, Xq=ops.quantize_acts(X))
                return ops.mul_mat_vec_q(t, Wm, X, N, K)
            for i in range(3):
                run(pool[i % len(pool)])
            torch.cuda.synchronize()
            # one CUDA graph of `iters` back-to-back launches: removes host launch latency from the
            # measurement (decode runs from a graph too); each launch still reads a different matrix
            iters = max(len(pool), 24)
            g = torch.cuda.CUDAGraph()
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                with torch.cuda.graph(g, stream=st):
                    for i in range(iters):
                        run(pool[i % len(pool)])
            ts = []
            for rep in range(6):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                g.replay()
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1) * 1e3 / iters)
            ts = sorted(ts[1:])
            med = ts[len(ts) // 2]
            gbs = wbytes / med / 1e3
            print("M=%d %-30s %8.2f us (min %7.2f)  %7.1f GB/s  %.3f of %s peak" % (M, name, med, ts[0], gbs, gbs / pk, how), flush=True)
            del pool


if __name__ == "__main__":
    main()
