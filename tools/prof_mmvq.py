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


def main():
    dev = torch.device("cuda:0")
    names = sys.argv[1:] or list(SHAPES)
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
