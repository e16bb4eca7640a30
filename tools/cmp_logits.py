"""cmp_logits.py a.bin b.bin -- NMSE between two logits_check outputs (test infrastructure)."""
import sys
import numpy as np


def load(p):
    raw = np.fromfile(p, dtype=np.uint8)
    nv = int(raw[:4].view(np.int32)[0])
    return raw[4:].view(np.float32).reshape(-1, nv)


a, b = load(sys.argv[1]), load(sys.argv[2])
worst = 0.0
for i in range(a.shape[0]):
    n = float(((a[i] - b[i]) ** 2).sum() / max((a[i] ** 2).sum(), 1e-30))
    worst = max(worst, n)
print("%s vs %s: %d steps, worst NMSE %.3e, identical=%s, argmax agree %d/%d" % (
    sys.argv[1], sys.argv[2], a.shape[0], worst, bool((a == b).all()), int((a.argmax(1) == b.argmax(1)).sum()), a.shape[0]))
