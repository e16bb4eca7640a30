"""ctypes front-end of the CPU oracle (oracle/oracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline
leg.  The product (llamacog_b200/) never imports this package.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "liboracle.so")
REF_LIB_PATH = os.path.join(HERE, "_ref", "libggml-ref.so")

# ggml type ids (reference ggml/include/ggml.h:352-366)
F32, F16, Q4_0, Q8_0, Q8_1, Q4_K, Q5_K, Q6_K = 0, 1, 2, 8, 9, 12, 13, 14
Q4_1, Q5_0, Q5_1 = 3, 6, 7
Q2_K, Q3_K = 10, 11
IQ4_NL, IQ4_XS = 20, 23
QUANT_TYPES = {"q4_0": Q4_0, "q8_0": Q8_0, "q4_K": Q4_K, "q5_K": Q5_K, "q6_K": Q6_K}
# the round-2 widening: the remaining 32-element block formats
LEGACY_TYPES = {"q4_1": Q4_1, "q5_0": Q5_0, "q5_1": Q5_1}
# ... and the two small K-quants (decode mat-vec, dequantisation and GET_ROWS; prefill runs as mat-vec passes)
SMALL_K_TYPES = {"q2_K": Q2_K, "q3_K": Q3_K, "iq4_nl": IQ4_NL, "iq4_xs": IQ4_XS}      # ... and the two 4-bit table formats
ALL_QUANT_TYPES = {**QUANT_TYPES, **LEGACY_TYPES, **SMALL_K_TYPES}
BLOCK_BYTES = {IQ4_NL: 18, IQ4_XS: 136, Q2_K: 84, Q3_K: 110, Q4_1: 20, Q5_0: 22, Q5_1: 24, Q4_0: 18, Q8_0: 34, Q8_1: 36, Q4_K: 144, Q5_K: 176, Q6_K: 210, F32: 4, F16: 2}
BLOCK_ELEMS = {IQ4_NL: 32, IQ4_XS: 256, Q2_K: 256, Q3_K: 256, Q4_1: 32, Q5_0: 32, Q5_1: 32, Q4_0: 32, Q8_0: 32, Q8_1: 32, Q4_K: 256, Q5_K: 256, Q6_K: 256, F32: 1, F16: 1}


def row_bytes(t, k):
    assert k % BLOCK_ELEMS[t] == 0
    return k // BLOCK_ELEMS[t] * BLOCK_BYTES[t]


def build(force=False):
    src = [os.path.join(HERE, f) for f in os.listdir(HERE) if f.endswith(".c")]
    stale = (not os.path.exists(LIB_PATH)) or any(os.path.getmtime(s) > os.path.getmtime(LIB_PATH) for s in src)
    if force or stale:
        subprocess.check_call(["make", "-s", "-C", HERE, "all"])


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(LIB_PATH)
        _lib.oracle_fp16_to_fp32.restype = C.c_float
        _lib.oracle_fp16_to_fp32.argtypes = [C.c_uint16]
        _lib.oracle_fp32_to_fp16.restype = C.c_uint16
        _lib.oracle_fp32_to_fp16.argtypes = [C.c_float]
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def dequantize_row(t, data, k):
    data = np.ascontiguousarray(data, dtype=np.uint8)
    y = np.empty(k, dtype=np.float32)
    rc = lib().oracle_dequantize_row(C.c_int(t), _p(data), _p(y), C.c_int64(k))
    assert rc == 0
    return y


def dequantize(t, data, n, k):
    rb = row_bytes(t, k)
    data = np.ascontiguousarray(data, dtype=np.uint8).reshape(n, rb)
    return np.stack([dequantize_row(t, data[i], k) for i in range(n)])


def quantize_row_q8_1_cuda(x, k_pad=None):
    x = np.ascontiguousarray(x, dtype=np.float32)
    k = x.size
    k_pad = k_pad or (k + 31) // 32 * 32
    y = np.zeros(k_pad // 32 * 36, dtype=np.uint8)
    lib().oracle_quantize_row_q8_1_cuda(_p(x), _p(y), C.c_int64(k), C.c_int64(k_pad))
    return y


def quantize_row_q8_0_ref(x):
    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.zeros(x.size // 32 * 34, dtype=np.uint8)
    lib().oracle_quantize_row_q8_0_ref(_p(x), _p(y), C.c_int64(x.size))
    return y


def quantize_row_q8_1_ref(x):
    x = np.ascontiguousarray(x, dtype=np.float32)
    y = np.zeros(x.size // 32 * 36, dtype=np.uint8)
    lib().oracle_quantize_row_q8_1_ref(_p(x), _p(y), C.c_int64(x.size))
    return y


def block_sums(t, wrow, q8_1, k):
    wrow = np.ascontiguousarray(wrow, dtype=np.uint8)
    q8_1 = np.ascontiguousarray(q8_1, dtype=np.uint8)
    si = np.zeros(k // 32, dtype=np.int32)
    sq = np.zeros(k // 32, dtype=np.int32)
    rc = lib().oracle_block_sums(C.c_int(t), _p(wrow), _p(q8_1), C.c_int64(k), _p(si), _p(sq))
    assert rc == 0
    return si, sq


def mul_mat_vec_q(t, W, X, N, K, M, use_s8=False, row_stride=None):
    """dst[M, N] (row m = activation column m), double-accumulated from exact int sums."""
    W = np.ascontiguousarray(W, dtype=np.uint8)
    X = np.ascontiguousarray(X, dtype=np.float32).reshape(M, K)
    dst = np.zeros((M, N), dtype=np.float32)
    rs = row_stride or row_bytes(t, K)
    rc = lib().oracle_mul_mat_vec_q(C.c_int(t), _p(W), C.c_int64(rs), _p(X), C.c_int64(K), _p(dst), C.c_int64(N),
                                    C.c_int64(N), C.c_int64(K), C.c_int64(M), C.c_int(int(use_s8)))
    assert rc == 0
    return dst


def mul_mat_f32(t, W, X, N, K, M):
    W = np.ascontiguousarray(W, dtype=np.uint8)
    X = np.ascontiguousarray(X, dtype=np.float32).reshape(M, K)
    dst = np.zeros((M, N), dtype=np.float32)
    rc = lib().oracle_mul_mat_f32(C.c_int(t), _p(W), C.c_int64(row_bytes(t, K)), _p(X), C.c_int64(K), _p(dst),
                                  C.c_int64(N), C.c_int64(N), C.c_int64(K), C.c_int64(M))
    assert rc == 0
    return dst


def random_quant_bytes(t, n, k, rng, scale=0.02):
    """Synthetic, well-formed quantised rows: random payload bits, sane finite fp16 scales.
    Exercises every code value (unlike a quantiser's output); used for parity and bench inputs."""
    rb = row_bytes(t, k)
    nb = n * (k // BLOCK_ELEMS[t])
    raw = rng.integers(0, 256, size=(nb, BLOCK_BYTES[t]), dtype=np.uint8)

    def put_half(col, vals):
        raw[:, col:col + 2] = vals.astype(np.float16).view(np.uint8).reshape(nb, 2)

    if t in (Q4_0, Q8_0, Q5_0):
        put_half(0, rng.uniform(0.2, 1.0, nb) * scale * rng.choice([-1.0, 1.0], nb))
    elif t in (Q4_1, Q5_1):
        put_half(0, rng.uniform(0.2, 1.0, nb) * scale)
        put_half(2, rng.uniform(-8.0, 8.0, nb) * scale)
    elif t in (Q4_K, Q5_K):
        put_half(0, rng.uniform(0.2, 1.0, nb) * scale / 32)
        put_half(2, rng.uniform(0.2, 1.0, nb) * scale / 32)
    elif t == Q6_K:
        put_half(208, rng.uniform(0.2, 1.0, nb) * scale / 64 * rng.choice([-1.0, 1.0], nb))
    elif t == Q2_K:
        put_half(80, rng.uniform(0.2, 1.0, nb) * scale / 16)
        put_half(82, rng.uniform(0.2, 1.0, nb) * scale / 16)
    elif t == IQ4_NL:
        put_half(0, rng.uniform(0.2, 1.0, nb) * scale / 64 * rng.choice([-1.0, 1.0], nb))
    elif t == IQ4_XS:
        put_half(0, rng.uniform(0.2, 1.0, nb) * scale / 1024 * rng.choice([-1.0, 1.0], nb))
    elif t == Q3_K:
        put_half(108, rng.uniform(0.2, 1.0, nb) * scale / 64 * rng.choice([-1.0, 1.0], nb))
    return raw.reshape(n, rb)
