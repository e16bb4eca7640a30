"""fa_trace.py -- phase timeline of the last query tile's CTA of the tensor-core attention kernel (library built with B200_KTRACE=1).
Marks (ns after the CTA's Q staging): 2 Q staged, 3 tile list ready, per KV tile: 4 S arrived, 5 S in registers, 6 softmax done,
7 previous P V finished, 8 P stored; 9 last P V finished, 10 output stored."""
import ctypes as C
import math
import sys

import torch

sys.path.insert(0, ".")
from llamacog_b200 import ops

N = int(sys.argv[1]) if len(sys.argv) > 1 else 512
n_kv = int(sys.argv[2]) if len(sys.argv) > 2 else N
D, n_head, n_head_kv = 128, 32, 8
q = torch.randn(1, n_head, N, D, device="cuda")
k = torch.randn(1, n_head_kv, n_kv, D, device="cuda").half()
v = torch.randn(1, n_head_kv, n_kv, D, device="cuda").half()
Np = (N + 63) // 64 * 64; Kp = (n_kv + 255) // 256 * 256
mask = torch.full((Np, Kp), float("-inf"), dtype=torch.float16, device="cuda")
for i in range(N):
    mask[i, : n_kv - N + i + 1] = 0
scale = 1.0 / math.sqrt(D)
L = ops.lib()
for _ in range(3):
    ops.flash_attn_ext(q, k, v, mask, scale)
torch.cuda.synchronize()
assert L.b200_ktrace_enable() == 0, "library built without -DB200_KTRACE"
for _ in range(3):
    ops.flash_attn_ext(q, k, v, mask, scale)
torch.cuda.synchronize()
path = "gpurun_out/fa_ktrace_%d_%d.txt" % (N, n_kv)
L.b200_ktrace_dump(path.encode())
ev = [tuple(int(x) for x in l.split()) for l in open(path)]
fa = [e for e in ev if e[4] == 192]
gids = sorted(set(e[1] for e in fa))
for gid in gids[-1:]:
    es = sorted(e for e in fa if e[1] == gid)
    t0 = es[0][0]
    print("grid %d (%d CTAs): " % (gid, es[0][3]) + " ".join("e%d@%.2f" % (e[2], (e[0] - t0) / 1e3) for e in es))
