#!/usr/bin/env python
"""q8_1 activation bytes made by the REFERENCE's own CUDA kernel (ggml-cuda/quantize.cu:4-49, quantize_row_q8_1_cuda), run on a B200.

The build container has no GPU, so quant_golden.npz's `q81cuda` is a numpy restatement of that kernel (make_golden.py).  This script
closes the gap: it loads baseline/_ref/lib/libggml-cuda.so (the unmodified reference backend built for sm_100 by `make -C baseline
cuda`), calls the exported quantize_row_q8_1_cuda on seeded inputs and writes tests/golden/q81_cuda_ref.npz:

    X      [rows, K] f32 inputs: uniform, normal, tiny / huge magnitudes, all-zero blocks, and values placed on the .5 rounding
           boundaries of x / d (where a reciprocal-multiply quantiser would differ from the reference's division)
    Y      [rows, K/32*36] u8: block_q8_1 {half d, half sum, int8 q[32]} as the reference kernel wrote them
    also the first two rows of quant_golden.npz's X (K = 512) as Y512, to cross-check the numpy-made `q81cuda`.

Run on the GPU box:  gpurun -- 'python tests/golden/make_q81_cuda_golden.py gpurun_out/q81_cuda_ref.npz'  then copy into tests/golden/.
"""
import ctypes as C
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
LIBDIR = os.path.join(ROOT, "baseline", "_ref", "lib")
SYM = "_Z22quantize_row_q8_1_cudaPKfPKiPv9ggml_typellllllllP11CUstream_st"
GGML_TYPE_Q4_K = 12


def ref_quantize(lib, x):
    """x: [rows, K] f32 on the host -> [rows, K/32*36] u8 from the reference kernel (K % 32 == 0, no padding columns)."""
    rows, k = x.shape
    xd = torch.from_numpy(x).cuda()
    yd = torch.zeros(rows * (k // 32) * 36, dtype=torch.uint8, device="cuda")
    fn = getattr(lib, SYM)
    fn.restype = None
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int] + [C.c_int64] * 8 + [C.c_void_p]
    # (x, ids, vy, type_src0, ne00, s01, s02, s03, ne0, ne1, ne2, ne3, stream): one column per row of X
    fn(xd.data_ptr(), None, yd.data_ptr(), GGML_TYPE_Q4_K, k, k, rows * k, rows * k, k, rows, 1, 1, None)
    torch.cuda.synchronize()
    return yd.cpu().numpy().reshape(rows, -1)


def boundary_rows(rng, k):
    """Blocks whose elements sit exactly on (j + 0.5) * d for the block's own d = amax / 127."""
    x = np.zeros(k, dtype=np.float32)
    for b in range(k // 32):
        amax = np.float32(rng.uniform(0.01, 100.0))
        d = np.float32(amax / np.float32(127))
        j = rng.integers(-126, 126, 32)
        v = ((j.astype(np.float64) + 0.5) * np.float64(d)).astype(np.float32)
        # nudge half of them by one ulp either way
        nud = rng.integers(-1, 2, 32)
        v = np.where(nud > 0, np.nextafter(v, np.float32(np.inf)), np.where(nud < 0, np.nextafter(v, np.float32(-np.inf)), v)).astype(np.float32)
        v[0] = amax * (1 if b % 2 == 0 else -1)
        x[32 * b: 32 * b + 32] = v
    return x


def main():
    out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "tests", "golden", "q81_cuda_ref.npz")
    C.CDLL(os.path.join(LIBDIR, "libggml-base.so"), mode=C.RTLD_GLOBAL)
    for dep in ("libggml-cpu.so", "libggml.so"):
        p = os.path.join(LIBDIR, dep)
        if os.path.exists(p):
            try:
                C.CDLL(p, mode=C.RTLD_GLOBAL)
            except OSError:
                pass
    lib = C.CDLL(os.path.join(LIBDIR, "libggml-cuda.so"), mode=C.RTLD_GLOBAL)
    rng = np.random.default_rng(20261105)
    k = 1024
    rows = [rng.uniform(-1, 1, k), rng.standard_normal(k) * 3, rng.standard_normal(k) * 1e-6, rng.standard_normal(k) * 1e4,
            rng.standard_normal(k) * 1e-30, rng.standard_normal(k) * 1e-41]
    z = rng.standard_normal(k)
    z[64:160] = 0.0
    rows.append(z)
    for _ in range(9):
        rows.append(boundary_rows(rng, k))
    x = np.stack(rows).astype(np.float32)
    y = ref_quantize(lib, x)
    g = np.load(os.path.join(ROOT, "tests", "golden", "quant_golden.npz"))
    x512 = np.ascontiguousarray(g["X"][:2]).astype(np.float32)
    y512 = ref_quantize(lib, x512)
    same = bool(np.array_equal(y512, g["q81cuda"][:2]))
    np.savez_compressed(out, X=x, Y=y, Y512=y512, device=np.array(torch.cuda.get_device_name(0)))
    print("wrote", out, x.shape, y.shape, "numpy-made q81cuda of quant_golden.npz == reference kernel:", same)


if __name__ == "__main__":
    main()
