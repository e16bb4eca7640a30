#!/usr/bin/env python
"""Tiny driver for ncu: a few direct launches of the prefill MMQ kernel on chosen shapes."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from llamacog_b200 import ops  # noqa: E402
from tools.mmvq_micro import rand_weights  # noqa: E402

SHAPES = {"up": (ops.Q4_K, 14336, 4096, 512), "q": (ops.Q4_K, 4096, 4096, 512), "down": (ops.Q4_K, 4096, 14336, 512),
          "up2k": (ops.Q4_K, 14336, 4096, 2048), "down6k": (ops.Q6_K, 4096, 14336, 512)}


def main():
    dev = torch.device("cuda:0")
    for n in sys.argv[1:] or ["up", "q"]:
        t, N, K, M = SHAPES[n]
        W = rand_weights(t, N * ops.row_bytes(t, K) + 256, dev)
        X = torch.randn((M, K), device=dev)
        for _ in range(3):
            ops.mul_mat_q(t, W, X, N, K)
        torch.cuda.synchronize()
    print("done")


if __name__ == "__main__":
    main()
