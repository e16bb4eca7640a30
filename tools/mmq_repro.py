#!/usr/bin/env python
"""Reproduce in-model MMQ shapes with really quantised weights (reference quantiser) against dequantise + f32 matmul."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
import oracle as O  # noqa: E402
from llamacog_b200 import ops  # noqa: E402
from make_gguf import Quantizer  # noqa: E402

q = Quantizer()
for name in sys.argv[1:] or ["q4_0", "q4_K", "q8_0"]:
    t = O.QUANT_TYPES[name]
    for (N, K, M) in [(256, 256, 64), (512, 256, 64), (256, 512, 64), (2048, 2048, 64), (512, 2048, 64), (8192, 2048, 64), (2048, 8192, 64),
                      (2048, 2048, 48), (2048, 2048, 33), (2048, 2048, 32), (2048, 2048, 100), (2048, 2048, 200)]:
        for kind in ("real", "random"):
            if kind == "real":
                W = q.tensor(t, N, K, 3)
            else:
                W = O.random_quant_bytes(t, N, K, np.random.default_rng(1))
            Wd = torch.zeros(W.size + 256, dtype=torch.uint8, device="cuda"); Wd[:W.size] = torch.from_numpy(W.reshape(-1)).cuda()
            X = torch.randn(M, K, device="cuda")
            got = ops.mul_mat_q(t, Wd, X, N, K)
            ref = X.double() @ torch.from_numpy(O.dequantize(t, W, N, K)).cuda().double().T
            nmse = float(((got.double() - ref) ** 2).sum() / (ref ** 2).sum())
            # per row-tile error to localise
            per_tile = [float(((got[:, i:i + 128].double() - ref[:, i:i + 128]) ** 2).sum() / (ref[:, i:i + 128] ** 2).sum()) for i in range(0, min(N, 512), 128)]
            print("%-5s %-6s N=%5d K=%5d M=%4d nmse %.3g  tiles %s" % (name, kind, N, K, M, nmse, " ".join("%.2g" % v for v in per_tile)), flush=True)
