"""fa_micro.py -- device time of FLASH_ATTN_EXT at long context (f16 and q8_0 KV), prefill and decode rows."""
import math
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from llamacog_b200 import ops


def bench(fn, n=5):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


D, n_head, n_head_kv = 128, 32, 8
for N, n_kv in [(512, 512), (512, 4096), (512, 33280), (1, 32768), (1, 4096), (8, 32768)]:
    n_past = n_kv - N
    q = torch.randn(1, n_head, N, D, device="cuda")
    k = torch.randn(1, n_head_kv, n_kv, D, device="cuda").half()
    v = torch.randn(1, n_head_kv, n_kv, D, device="cuda").half()
    Np = (N + 63) // 64 * 64; Kp = (n_kv + 255) // 256 * 256
    mask = torch.full((Np, Kp), float("-inf"), dtype=torch.float16, device="cuda")
    for i in range(N):
        mask[i, : n_past + i + 1] = 0
    k8 = torch.randint(0, 255, (1, n_head_kv, n_kv, D // 32 * 34), dtype=torch.uint8, device="cuda")
    k8.view(-1, 34)[:, 1] = 0x20            # sane fp16 scales
    v8 = k8.clone()
    scale = 1.0 / math.sqrt(D)
    t16 = bench(lambda: ops.flash_attn_ext(q, k, v, mask, scale))
    t8 = bench(lambda: ops.flash_attn_ext(q, k8, v8, mask, scale))
    flops = 4.0 * N * (n_past + (N + 1) / 2) * D * n_head
    kvb16 = 2 * n_kv * n_head_kv * D * 2; kvb8 = kvb16 * 34 / 64
    print("N=%d n_kv=%d: f16 %.3f ms (%.1f TF/s, %.0f GB/s KV)   q8_0 %.3f ms (%.1f TF/s, %.0f GB/s KV)" % (
        N, n_kv, t16, flops / t16 / 1e9, kvb16 / t16 / 1e6, t8, flops / t8 / 1e9, kvb8 / t8 / 1e6))
