import sys, numpy as np
a = np.fromfile(sys.argv[1], dtype=np.float32); b = np.fromfile(sys.argv[2], dtype=np.float32)
n = min(a.size, b.size); d = np.abs(a[:n] - b[:n]); i = int(d.argmax())
V = 128256
rows = sorted(set((np.nonzero(d)[0] // V).tolist()))
print("sizes %d %d  differing values %d  max |diff| %g at row %d col %d (values %g %g)  rows with differences: %s" % (a.size, b.size, int((d != 0).sum()), d[i], i // V, i % V, a[i], b[i], rows[:12]))
