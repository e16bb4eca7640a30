#!/usr/bin/env python
"""Generate tests/golden/quant_golden.npz from the REFERENCE's own code.

Run in the build container only (needs oracle/_ref/libggml-ref.so, which oracle/Makefile compiles
from /root/reference/ggml/src where the sources lie).  The committed .npz travels to the GPU box,
where /root/reference does not exist.

What is pinned, and by which reference function:
  * W_<t> / deq_<t>      ggml_quantize_chunk + dequantize_row_<t>          (ggml-quants.c)
  * R_<t> / rdeq_<t>     random well-formed block bytes through dequantize_row_<t>
  * q80ref / q81ref      quantize_row_q8_0_ref / quantize_row_q8_1_ref     (ggml-quants.c:187,213)
  * h2f / f2h            ggml_fp16_to_fp32 (all 65536 codes) / ggml_fp32_to_fp16
  * sumi_<t> / sumq      int32 block dot products of ggml-cuda/vecdotq.cuh, derived here by an
                         INDEPENDENT route: integer codes recovered from the reference's
                         dequantised floats (not from our bit-unpacking), dotted in numpy with
  * q81cuda              a numpy restatement of ggml-cuda/quantize.cu:4-49 (the CUDA kernel itself
                         cannot run in the container: no GPU).  make_q81_cuda_golden.py ran the reference's
                         kernel on a B200 afterwards: its bytes for these rows equal q81cuda (Y512 of
                         q81_cuda_ref.npz, checked by tests/test_oracle_golden.py).
"""
import ctypes as C
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import oracle as O  # noqa: E402  (only for type ids / sizes / random_quant_bytes)

ref = C.CDLL(O.REF_LIB_PATH)
ref.ggml_fp16_to_fp32.restype = C.c_float
ref.ggml_fp16_to_fp32.argtypes = [C.c_uint16]
ref.ggml_fp32_to_fp16.restype = C.c_uint16
ref.ggml_fp32_to_fp16.argtypes = [C.c_float]
ref.ggml_quantize_chunk.restype = C.c_size_t
ref.ggml_quantize_chunk.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]



class _InitParams(C.Structure):
    _fields_ = [("mem_size", C.c_size_t), ("mem_buffer", C.c_void_p), ("no_alloc", C.c_bool)]


# ggml_init fills the fp16->fp32 lookup table every dequantiser reads (ggml.c:1419-1433)
ref.ggml_init.restype = C.c_void_p
ref.ggml_init.argtypes = [_InitParams]
assert ref.ggml_init(_InitParams(1 << 20, None, False))

K, N, M = 512, 4, 2


def p(a):
    return a.ctypes.data_as(C.c_void_p)


def ref_dequant(name, data, n, k):
    fn = getattr(ref, "dequantize_row_" + name)
    out = np.empty((n, k), dtype=np.float32)
    rb = O.row_bytes(O.ALL_QUANT_TYPES[name], k)
    data = np.ascontiguousarray(data).reshape(n, rb)
    for i in range(n):
        row = np.ascontiguousarray(data[i])
        fn(p(row), p(out[i]), C.c_int64(k))
    return out


def roundf(v):  # C roundf: half away from zero
    v = v.astype(np.float64)
    return np.sign(v) * np.floor(np.abs(v) + 0.5)


def q8_1_cuda_numpy(x, k_pad):
    """numpy restatement of quantize.cu:4-49 for one row -> (bytes, q int8 [k_pad], d f32 [k_pad/32])"""
    xp = np.zeros(k_pad, dtype=np.float32)
    xp[: x.size] = x
    b = xp.reshape(-1, 32)
    amax = np.abs(b).max(axis=1)
    d = (amax / np.float32(127)).astype(np.float32)
    with np.errstate(divide="ignore", invalid="ignore"):
        q = np.where(amax[:, None] == 0, 0, roundf((b / d[:, None]).astype(np.float32))).astype(np.int8)
    s = b.copy()
    for off in (16, 8, 4, 2, 1):
        s = (s + s[:, np.arange(32) ^ off]).astype(np.float32)
    out = np.zeros((b.shape[0], 36), dtype=np.uint8)
    out[:, 0:2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 2:4] = s[:, 0].astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 4:] = q.view(np.uint8)
    return out.reshape(-1), q.reshape(-1), d


def int_codes_from_dequant(name, data, deq, n, k):
    """Recover the integer code of every weight from the reference's dequantised floats."""
    t = O.ALL_QUANT_TYPES[name]
    rb = O.row_bytes(t, k)
    raw = np.ascontiguousarray(data).reshape(n, rb)
    codes = np.zeros((n, k), dtype=np.int64)
    for i in range(n):
        blocks = raw[i].reshape(-1, O.BLOCK_BYTES[t])
        for b, blk in enumerate(blocks):
            if name in ("q4_0", "q8_0", "q5_0"):
                d = np.float64(blk[0:2].view(np.float16)[0])
                y = deq[i, 32 * b: 32 * b + 32].astype(np.float64)
                c = np.rint(y / d) + {"q4_0": 8, "q8_0": 0, "q5_0": 16}[name]
                codes[i, 32 * b: 32 * b + 32] = c
            elif name in ("q4_1", "q5_1"):
                d = np.float64(blk[0:2].view(np.float16)[0])
                m = np.float64(blk[2:4].view(np.float16)[0])
                assert d != 0
                y = deq[i, 32 * b: 32 * b + 32].astype(np.float64)
                codes[i, 32 * b: 32 * b + 32] = np.rint((y - m) / d)
            elif name in ("q4_K", "q5_K"):
                d = np.float32(blk[0:2].view(np.float16)[0])
                dm = np.float32(blk[2:4].view(np.float16)[0])
                sc6 = blk[4:16]
                for j in range(8):
                    if j < 4:
                        sc, mn = sc6[j] & 63, sc6[j + 4] & 63
                    else:
                        sc = (sc6[j + 4] & 0xF) | ((sc6[j - 4] >> 6) << 4)
                        mn = (sc6[j + 4] >> 4) | ((sc6[j] >> 6) << 4)
                    y = deq[i, 256 * b + 32 * j: 256 * b + 32 * j + 32].astype(np.float64)
                    ds = np.float64(np.float32(d * np.float32(sc)))
                    ms = np.float64(np.float32(dm * np.float32(mn)))
                    assert sc != 0, "golden weights must have non-zero sub-scales for code recovery"
                    codes[i, 256 * b + 32 * j: 256 * b + 32 * j + 32] = np.rint((y + ms) / ds)
            elif name == "iq4_nl":
                d = np.float64(blk[0:2].view(np.float16)[0])
                assert d != 0
                codes[i, 32 * b: 32 * b + 32] = np.rint(deq[i, 32 * b: 32 * b + 32].astype(np.float64) / d)
            elif name == "iq4_xs":
                d = np.float32(blk[0:2].view(np.float16)[0]); sh = int(blk[2:4].view(np.uint16)[0]); sl = blk[4:8]
                for ib in range(8):
                    ls = (((int(sl[ib // 2]) >> (4 * (ib % 2))) & 0xF) | (((sh >> (2 * ib)) & 3) << 4)) - 32
                    assert ls != 0 and d != 0, "golden weights must have non-zero block scales for code recovery"
                    dl = np.float64(np.float32(d * np.float32(ls)))
                    codes[i, 256 * b + 32 * ib: 256 * b + 32 * ib + 32] = np.rint(deq[i, 256 * b + 32 * ib: 256 * b + 32 * ib + 32].astype(np.float64) / dl)
            elif name == "q2_K":
                d = np.float32(blk[80:82].view(np.float16)[0]); dm = np.float32(blk[82:84].view(np.float16)[0])
                for g in range(16):
                    sc = blk[g]
                    assert (sc & 0xF) != 0, "golden weights must have non-zero sub-scales for code recovery"
                    dl = np.float64(np.float32(d * np.float32(sc & 0xF))); ml = np.float64(np.float32(dm * np.float32(sc >> 4)))
                    y = deq[i, 256 * b + 16 * g: 256 * b + 16 * g + 16].astype(np.float64)
                    codes[i, 256 * b + 16 * g: 256 * b + 16 * g + 16] = np.rint((y + ml) / dl)
            elif name == "q3_K":
                d = np.float32(blk[108:110].view(np.float16)[0])
                sc12 = blk[96:108]
                for g in range(16):
                    bb, gg = g & 3, g >> 2
                    lo = (sc12[4 * gg + bb] & 0xF) if gg < 2 else (sc12[4 * (gg - 2) + bb] >> 4)
                    scv = int(lo | (((sc12[8 + bb] >> (2 * gg)) & 3) << 4)) - 32
                    assert scv != 0, "golden weights must have non-zero sub-scales for code recovery"
                    dl = np.float64(np.float32(d * np.float32(scv)))
                    y = deq[i, 256 * b + 16 * g: 256 * b + 16 * g + 16].astype(np.float64)
                    codes[i, 256 * b + 16 * g: 256 * b + 16 * g + 16] = np.rint(y / dl)
            elif name == "q6_K":
                d = np.float32(blk[208:210].view(np.float16)[0])
                scs = blk[192:208].view(np.int8)
                for g in range(16):
                    assert scs[g] != 0
                    ds = np.float64(np.float32(d * np.float32(scs[g])))
                    y = deq[i, 256 * b + 16 * g: 256 * b + 16 * g + 16].astype(np.float64)
                    codes[i, 256 * b + 16 * g: 256 * b + 16 * g + 16] = np.rint(y / ds)
    return codes


def main():
    rng = np.random.default_rng(20261104)
    out = {}
    # fp16 tables
    out["h2f"] = np.array([ref.ggml_fp16_to_fp32(i) for i in range(65536)], dtype=np.float32).view(np.uint32)
    fl = np.concatenate([
        rng.standard_normal(2048).astype(np.float32) * np.float32(10.0) ** rng.integers(-8, 5, 2048).astype(np.float32),
        np.array([0.0, -0.0, 65504.0, 65519.9, 65520.0, 1e-8, 5.96e-8, 2.98e-8, 2.9802322e-8, 6.1e-5, 6.0975552e-5,
                  np.inf, -np.inf, 1.0009765625, 1.00048828125, 1.000732421875], dtype=np.float32)])
    out["f2h_in"] = fl
    out["f2h"] = np.array([ref.ggml_fp32_to_fp16(float(v)) for v in fl], dtype=np.uint16)

    # activations: uniform(-1,1) like tests/test-backend-ops.cpp:46-53, plus a zero block and an outlier
    X = rng.uniform(-1, 1, size=(M, K)).astype(np.float32)
    X[0, 64:96] = 0.0
    X[1, 200] = 37.5
    out["X"] = X
    q80 = np.zeros((M, K // 32 * 34), dtype=np.uint8)
    q81 = np.zeros((M, K // 32 * 36), dtype=np.uint8)
    for m in range(M):
        ref.quantize_row_q8_0_ref(p(X[m]), p(q80[m]), C.c_int64(K))
        ref.quantize_row_q8_1_ref(p(X[m]), p(q81[m]), C.c_int64(K))
    out["q80ref"], out["q81ref"] = q80, q81
    k_pad = 512
    cu = [q8_1_cuda_numpy(X[m], k_pad) for m in range(M)]
    out["q81cuda"] = np.stack([c[0] for c in cu])
    qx = np.stack([c[1] for c in cu]).astype(np.int64)  # [M, K]
    out["sumq"] = qx.reshape(M, K // 32, 32).sum(axis=2).astype(np.int32)

    Wf = (rng.standard_normal((N, K)) * 0.02).astype(np.float32)
    out["Wf"] = Wf
    for name, t in O.QUANT_TYPES.items():
        rb = O.row_bytes(t, K)
        data = np.zeros(N * rb, dtype=np.uint8)
        sz = ref.ggml_quantize_chunk(t, p(Wf), p(data), 0, N, K, None)
        assert sz == N * rb
        deq = ref_dequant(name, data, N, K)
        out["W_" + name], out["deq_" + name] = data.reshape(N, rb), deq.view(np.uint32)
        rnd = O.random_quant_bytes(t, N, K, rng)
        rdeq = ref_dequant(name, rnd, N, K)
        out["R_" + name], out["rdeq_" + name] = rnd, rdeq.view(np.uint32)
        codes = int_codes_from_dequant(name, data, deq, N, K)
        # q6_K codes recovered from floats are already (q-32); q4_0 recovered +8 -> raw nibble
        sumi = np.einsum("nbk,mbk->mnb", codes.reshape(N, K // 32, 32), qx.reshape(M, K // 32, 32)).astype(np.int32)
        out["sumi_" + name] = sumi
    np.savez_compressed(os.path.join(HERE, "quant_golden.npz"), **out)
    print("wrote quant_golden.npz:", {k: v.shape for k, v in out.items()})


def main_legacy():
    """quant_golden_legacy.npz: the same vectors for q4_1 / q5_0 / q5_1 and q2_K / q3_K, over the activations X of quant_golden.npz so
    that q81cuda / sumq are shared.  A file of its own: quant_golden.npz stays byte-identical."""
    G = np.load(os.path.join(HERE, "quant_golden.npz"))
    X = G["X"]
    qx = np.stack([q8_1_cuda_numpy(X[m], 512)[1] for m in range(M)]).astype(np.int64)
    rng = np.random.default_rng(20261105)
    out = {}
    Wf = G["Wf"]
    for name, t in {**O.LEGACY_TYPES, **O.SMALL_K_TYPES}.items():
        rb = O.row_bytes(t, K)
        data = np.zeros(N * rb, dtype=np.uint8)
        sz = ref.ggml_quantize_chunk(t, p(Wf), p(data), 0, N, K, None)
        assert sz == N * rb
        deq = ref_dequant(name, data, N, K)
        out["W_" + name], out["deq_" + name] = data.reshape(N, rb), deq.view(np.uint32)
        rnd = O.random_quant_bytes(t, N, K, rng)
        rdeq = ref_dequant(name, rnd, N, K)
        out["R_" + name], out["rdeq_" + name] = rnd, rdeq.view(np.uint32)
        codes = int_codes_from_dequant(name, data, deq, N, K)
        lo, hi = {"q4_1": (0, 15), "q5_0": (0, 31), "q5_1": (0, 31), "q2_K": (0, 3), "q3_K": (-4, 3), "iq4_nl": (-127, 113), "iq4_xs": (-127, 113)}[name]
        assert codes.min() >= lo and codes.max() <= hi
        out["sumi_" + name] = np.einsum("nbk,mbk->mnb", codes.reshape(N, K // 32, 32), qx.reshape(M, K // 32, 32)).astype(np.int32)
    np.savez_compressed(os.path.join(HERE, "quant_golden_legacy.npz"), **out)
    print("wrote quant_golden_legacy.npz:", {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "legacy":
        main_legacy()
    else:
        main()
