"""fa_decode_micro.py -- device time and cache bandwidth of decode FLASH_ATTN_EXT (1 row) over n_kv, per kernel family.

Run once per setting (the switches are read once per process):  B200_FATTN_MMA=0|1 python tools/fa_decode_micro.py [f16|q8_0]
Prints us per call, GB/s of K + V bytes and the maximum deviation from the fp32 torch reference."""
import math
import os
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from llamacog_b200 import ops

kv = sys.argv[1] if len(sys.argv) > 1 else "f16"
only = int(sys.argv[2]) if len(sys.argv) > 2 else 0      # one context length, 32/8 heads (for ncu)
D = 128


def bench(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


print("B200_FATTN_MMA=%s kv=%s" % (os.environ.get("B200_FATTN_MMA", "(default)"), kv))
for n_head, n_head_kv in (((32, 8),) if only else ((32, 8), (64, 8))):
    for n_kv in ((only,) if only else (512, 1024, 2048, 4096, 8192, 16384, 32768)):
        # several caches so that consecutive calls do not hit L2 (8 heads x 32k x 512 B = 134 MB per cache)
        ncache = max(2, int(400e6 // (n_head_kv * n_kv * D * 4)) + 1)
        q = torch.randn(1, n_head, 1, D, device="cuda")
        ks = [torch.randn(1, n_head_kv, n_kv, D, device="cuda").half() for _ in range(ncache)]
        vs = [torch.randn(1, n_head_kv, n_kv, D, device="cuda").half() for _ in range(ncache)]
        Kp = (n_kv + 255) // 256 * 256
        mask = torch.zeros((64, Kp), dtype=torch.float16, device="cuda")
        mask[:, n_kv:] = float("-inf")
        scale = 1.0 / math.sqrt(D)
        if kv == "q8_0":
            def quant(x):      # block_q8_0 rows {half d; int8 q[32]} (timing only: plain round-to-nearest)
                b = x.float().reshape(-1, 32)
                d = (b.abs().amax(1, keepdim=True) / 127.0).half()
                qv = torch.where(d.float() > 0, torch.round(b / d.float()), torch.zeros_like(b)).clamp(-127, 127).to(torch.int8)
                out = torch.empty((b.shape[0], 34), dtype=torch.uint8, device=x.device)
                out[:, :2] = d.view(torch.uint8).reshape(-1, 2)
                out[:, 2:] = qv.view(torch.uint8)
                return out.reshape(1, n_head_kv, n_kv, (D // 32) * 34)
            kq = [quant(x) for x in ks]; vq = [quant(x) for x in vs]
            calls = [lambda i=i: ops.flash_attn_ext(q, kq[i], vq[i], mask, scale) for i in range(ncache)]
            nbytes = n_head_kv * n_kv * (D // 32) * 34 * 2
        else:
            calls = [lambda i=i: ops.flash_attn_ext(q, ks[i], vs[i], mask, scale) for i in range(ncache)]
            nbytes = n_head_kv * n_kv * D * 2 * 2
        out = calls[0]()
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()          # one graph over all the caches: no host time between the launches
        with torch.cuda.graph(g):
            for c in calls:
                c()
        ms = bench(g.replay, n=10) / len(calls)
        kf, vf = ks[0].float().repeat_interleave(n_head // n_head_kv, 1), vs[0].float().repeat_interleave(n_head // n_head_kv, 1)
        ref = torch.softmax((q @ kf.transpose(-1, -2)) * scale, -1) @ vf
        err = (out.reshape(ref.shape) - ref).abs().max().item() if kv == "f16" else float("nan")
        print("heads %2d/%d n_kv %6d: %8.2f us  %7.1f GB/s  max|err| %.2e" % (n_head, n_head_kv, n_kv, ms * 1e3, nbytes / ms / 1e6, err))
