import sys, numpy as np
ref = np.fromfile(sys.argv[1], dtype=np.float32)
V = 128256
for f in sys.argv[2:]:
    x = np.fromfile(f, dtype=np.float32)
    n = min(ref.size, x.size) // V
    out = []
    for r in range(n):
        a, b = ref[r * V:(r + 1) * V].astype(np.float64), x[r * V:(r + 1) * V].astype(np.float64)
        out.append("%.2e" % (((a - b) ** 2).sum() / (a ** 2).sum()))
    print(f, "NMSE per row vs reference:", " ".join(out))
