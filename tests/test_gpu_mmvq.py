"""GPU parity of the decode GEMV (through the C ABI) against the CPU oracle.

Bars: int32 block dot products bit-exact; fp32 outputs within 2e-5 of the oracle's double
accumulation relative to the output RMS (fp32 summation-order noise only); NMSE against the
dequantise-then-f32 matmul <= 5e-4 (reference tests/test-backend-ops.cpp:1990-1992)."""
import os

import numpy as np
import pytest
import torch

import oracle as O

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "quant_golden.npz"))
NAMES = list(O.QUANT_TYPES)


def _ops():
    from llamacog_b200 import ops
    return ops


def _run(t, Wnp, Xnp, N, K, dbg=False):
    ops = _ops()
    W = torch.from_numpy(np.ascontiguousarray(Wnp).reshape(-1)).cuda()
    X = torch.from_numpy(np.ascontiguousarray(Xnp)).cuda()
    out = ops.mul_mat_vec_q(t, W, X, N, K, dbg=dbg)
    torch.cuda.synchronize()
    return out


def _close(got, want, tol=2e-5):
    rms = float(np.sqrt((want.astype(np.float64) ** 2).mean())) + 1e-30
    err = float(np.abs(got.astype(np.float64) - want).max()) / rms
    assert err < tol, "max err / rms = %g" % err


@pytest.mark.parametrize("name", NAMES)
def test_golden_block_sums_bit_exact(name):
    """The reference-derived golden int32 sums (tests/golden) out of the production kernel."""
    t = O.QUANT_TYPES[name]
    dst, sums = _run(t, G["W_" + name], G["X"], 4, 512, dbg=True)
    assert np.array_equal(sums.cpu().numpy(), G["sumi_" + name])
    _close(dst.cpu().numpy(), O.mul_mat_vec_q(t, G["W_" + name], G["X"], 4, 512, 2))


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("M", [1, 2, 3, 4, 5, 8])
def test_small_shapes_vs_oracle(name, M):
    """m=16,k=256 n=1..8: the reference's own MUL_MAT grid (tests/test-backend-ops.cpp:4246-4250)."""
    t = O.QUANT_TYPES[name]
    rng = np.random.default_rng(100 + M)
    N, K = 16, 256
    W = O.random_quant_bytes(t, N, K, rng)
    X = rng.uniform(-1, 1, (M, K)).astype(np.float32)
    dst, sums = _run(t, W, X, N, K, dbg=True)
    for m in range(M):
        q = O.quantize_row_q8_1_cuda(X[m], K)
        for n in range(N):
            si, _ = O.block_sums(t, W[n], q, K)
            assert np.array_equal(sums[m, n].cpu().numpy(), si)
    _close(dst.cpu().numpy(), O.mul_mat_vec_q(t, W, X, N, K, M))


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("N,K,M", [(257, 4096, 1), (130, 14336, 1), (96, 4096, 4), (64, 2048, 8), (40, 1024, 7),
                                   (33, 8192, 2), (1, 512, 1), (7, 28672, 1)])
def test_model_shapes_vs_oracle(name, N, K, M):
    t = O.QUANT_TYPES[name]
    rng = np.random.default_rng(N * 7 + K + M)
    W = O.random_quant_bytes(t, N, K, rng)
    X = rng.uniform(-1, 1, (M, K)).astype(np.float32)
    X[0, :32] = 0.0                      # an all-zero q8_1 block (d == 0 path)
    dst = _run(t, W, X, N, K).cpu().numpy()
    want = O.mul_mat_vec_q(t, W, X, N, K, M)
    _close(dst, want)
    f32 = O.mul_mat_f32(t, W, X, N, K, M)
    nmse = float(((dst - f32) ** 2).sum() / (f32 ** 2).sum())
    assert nmse < 5e-4, nmse


@pytest.mark.parametrize("name", NAMES)
def test_strided_unaligned_rows(name):
    """Rows of a non-contiguous view (row stride > row bytes, 2-byte-aligned start), the layout
    test-backend-ops exercises with v=1 (tests/test-backend-ops.cpp:2043-2055)."""
    ops = _ops()
    t = O.QUANT_TYPES[name]
    rng = np.random.default_rng(5)
    N, K, M = 19, 512, 2
    rb = O.row_bytes(t, K)
    nb1 = rb + O.BLOCK_BYTES[t] * 3
    off = O.BLOCK_BYTES[t]              # start one block in: only 2-byte aligned for 18/34/210-byte blocks
    Wrows = O.random_quant_bytes(t, N, K, rng)
    buf = np.zeros(off + N * nb1 + 64, dtype=np.uint8)
    for n in range(N):
        buf[off + n * nb1: off + n * nb1 + rb] = Wrows[n]
    X = rng.uniform(-1, 1, (M, K)).astype(np.float32)
    Wd = torch.from_numpy(buf).cuda()
    dst = ops.mul_mat_vec_q(t, Wd[off:], torch.from_numpy(X).cuda(), N, K, nb1=nb1)
    torch.cuda.synchronize()
    _close(dst.cpu().numpy(), O.mul_mat_vec_q(t, Wrows, X, N, K, M))


@pytest.mark.parametrize("name", ["q4_K", "q6_K"])
def test_full_size_linearity(name):
    """BASELINE sizes (ffn_down 4096x14336): too big for the scalar oracle in seconds, so check a
    size-independent property -- the GEMV is linear in the weights' row order: permuting rows
    permutes outputs exactly (bit-for-bit), and 64 sampled rows match the oracle."""
    ops = _ops()
    t = O.QUANT_TYPES[name]
    rng = np.random.default_rng(11)
    N, K = 4096, 14336
    W = O.random_quant_bytes(t, N, K, rng)
    X = rng.uniform(-1, 1, (1, K)).astype(np.float32)
    Xd = torch.from_numpy(X).cuda()
    d0 = ops.mul_mat_vec_q(t, torch.from_numpy(W.reshape(-1)).cuda(), Xd, N, K).cpu().numpy()
    perm = rng.permutation(N)
    d1 = ops.mul_mat_vec_q(t, torch.from_numpy(np.ascontiguousarray(W[perm]).reshape(-1)).cuda(), Xd, N, K).cpu().numpy()
    assert np.array_equal(d1[0], d0[0][perm])
    rows = rng.choice(N, 64, replace=False)
    want = O.mul_mat_vec_q(t, np.ascontiguousarray(W[rows]), X, 64, K, 1)
    _close(d0[:, rows], want)


def test_quantize_q8_1_bit_exact():
    ops = _ops()
    rng = np.random.default_rng(3)
    X = rng.uniform(-1, 1, (5, 200)).astype(np.float32)
    X[2, 32:64] = 0
    X[3, 7] = 1e4
    y = ops.quantize_q8_1(torch.from_numpy(X).cuda(), 512).cpu().numpy()
    for r in range(5):
        assert np.array_equal(y[r], O.quantize_row_q8_1_cuda(X[r], 512))
    for m in range(2):
        y = ops.quantize_q8_1(torch.from_numpy(G["X"][m:m + 1].copy()).cuda(), 512).cpu().numpy()
        assert np.array_equal(y[0], G["q81cuda"][m])


def test_quantize_q8_1_vs_reference_kernel_bytes():
    """The product quantiser against bytes written by the reference's quantize_q8_1 kernel on a B200 (q81_cuda_ref.npz)."""
    from test_oracle_golden import _q81_ref_kernel_check
    ops = _ops()
    _q81_ref_kernel_check(lambda x: ops.quantize_q8_1(torch.from_numpy(x[None].copy()).cuda(), x.size).cpu().numpy()[0])


@pytest.mark.parametrize("name", NAMES)
def test_dequantize_bit_exact(name):
    ops = _ops()
    t = O.QUANT_TYPES[name]
    for src, ref in (("W_", "deq_"), ("R_", "rdeq_")):
        got = ops.dequantize_rows(t, torch.from_numpy(G[src + name].reshape(-1).copy()).cuda(), 4, 512).cpu().numpy()
        assert np.array_equal(got.view(np.uint32), G[ref + name])
    rng = np.random.default_rng(9)
    raw = O.random_quant_bytes(t, 8, 4096, rng)
    got = ops.dequantize_rows(t, torch.from_numpy(raw.reshape(-1).copy()).cuda(), 8, 4096).cpu().numpy()
    assert np.array_equal(got.view(np.uint32), O.dequantize(t, raw, 8, 4096).view(np.uint32))


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("N,K,M", [(64, 4096, 1), (37, 14336, 2), (16, 512, 4)])
def test_prequantised_path_matches_fused(name, N, K, M):
    """b200_quantize_acts + Xq (pre-quantised activations) must give the same int32 block sums and the
    same fp32 result, bit for bit, as the fused prologue: both run the same quantiser."""
    ops = _ops()
    t = O.QUANT_TYPES[name]
    rng = np.random.default_rng(K + M)
    W = torch.from_numpy(O.random_quant_bytes(t, N, K, rng).reshape(-1)).cuda()
    X = torch.from_numpy(rng.uniform(-1, 1, (M, K)).astype(np.float32)).cuda()
    d0, s0 = ops.mul_mat_vec_q(t, W, X, N, K, dbg=True)
    d1, s1 = ops.mul_mat_vec_q(t, W, X, N, K, dbg=True, Xq=ops.quantize_acts(X))
    torch.cuda.synchronize()
    assert torch.equal(s0, s1)
    if M == 1 and name in ("q4_K", "q6_K", "q8_0", "q4_0", "q5_K"):
        # batch-1 GEMVs of these formats on pre-quantised activations run the register-resident kernel (mmvq_r.cuh): same
        # integer sums, its own (fixed) fp32 summation order
        _close(d1.cpu().numpy(), d0.cpu().numpy().astype(np.float64))
    else:
        assert torch.equal(d0, d1)


RR_SHAPES = [(257, 4096), (130, 14336), (33, 8192), (7, 28672), (40, 1024), (64, 2048), (1, 512), (19, 5120), (16, 256),
             (1500, 4096), (3, 2560)]


@pytest.mark.parametrize("name", ["q4_K", "q6_K", "q8_0", "q4_0", "q5_K"])
@pytest.mark.parametrize("N,K", RR_SHAPES)
def test_register_resident_kernel_vs_oracle(name, N, K):
    """mmvq_r.cuh (M = 1, pre-quantised activations; one thread per 128-weight half item with its activations in
    registers): every int32 block sum bit-exact against the oracle, fp32 within summation-order noise.  The shapes cover
    threads-per-row 2..256, idle lanes (K = 5120, 14336, 28672), several rows per warp (K <= 2048), CTAs with a single
    row and multi-stage CTAs (N = 1500)."""
    ops = _ops()
    t = O.QUANT_TYPES[name]
    rng = np.random.default_rng(N * 31 + K)
    Wnp = O.random_quant_bytes(t, N, K, rng)
    Xnp = rng.uniform(-1, 1, (1, K)).astype(np.float32)
    W = torch.from_numpy(Wnp.reshape(-1)).cuda()
    X = torch.from_numpy(Xnp).cuda()
    dst, sums = ops.mul_mat_vec_q(t, W, X, N, K, dbg=True, Xq=ops.quantize_acts(X))
    dst2 = ops.mul_mat_vec_q(t, W, X, N, K, Xq=ops.quantize_acts(X))
    torch.cuda.synchronize()
    q = O.quantize_row_q8_1_cuda(Xnp[0], K)
    sums = sums.cpu().numpy()
    for n in range(0, N, max(1, N // 40)):
        si, _ = O.block_sums(t, Wnp[n], q, K)
        assert np.array_equal(sums[0, n], si), "row %d" % n
    want = O.mul_mat_vec_q(t, Wnp, Xnp, N, K, 1)
    _close(dst.cpu().numpy(), want)
    assert torch.equal(dst, dst2)          # parity-hook build == production build of the kernel


@pytest.mark.parametrize("name", ["q4_K", "q6_K", "q8_0", "q4_0", "q5_K"])
@pytest.mark.parametrize("Ns,K", [((4096, 1024, 1024), 4096), ((1432, 1432), 4096), ((5, 300), 14336), ((1, 1, 1), 512)])
def test_group_launch_matches_single(name, Ns, K):
    """q/k/v and gate/up share one launch (b200_mul_mat_vec_q_group): every matrix must come out bit-identical to its own
    single launch -- a row's fp32 sum does not depend on how rows are dealt to CTAs."""
    ops = _ops()
    t = O.QUANT_TYPES[name]
    rng = np.random.default_rng(sum(Ns) + K)
    Ws = [torch.from_numpy(O.random_quant_bytes(t, n, K, rng).reshape(-1)).cuda() for n in Ns]
    X = torch.from_numpy(rng.uniform(-1, 1, (1, K)).astype(np.float32)).cuda()
    Xq = ops.quantize_acts(X)
    got = ops.mul_mat_vec_q_group(t, Ws, list(Ns), X, K, Xq=Xq)
    for w, n, g in zip(Ws, Ns, got):
        one = ops.mul_mat_vec_q(t, w, X, n, K, Xq=Xq)
        torch.cuda.synchronize()
        assert torch.equal(one, g)


def test_fused_quantiser_rounding_boundaries():
    """The prologue replaces x/d by x*rcp(d) away from rounding boundaries; values placed on and next
    to every k+0.5 boundary (where the shortcut must fall back to the exact division) have to
    reproduce roundf(x/d) of the reference quantiser (quantize.cu:38-47) exactly."""
    ops = _ops()
    rng = np.random.default_rng(17)
    K = 4096
    X = np.empty((8, K), dtype=np.float32)
    for m in range(8):
        for b in range(K // 32):
            amax = np.float32(rng.uniform(0.01, 50.0))
            d = np.float32(amax / np.float32(127.0))
            k = rng.integers(-126, 126, 32).astype(np.float32) + np.float32(0.5)
            x = (k * d).astype(np.float32)
            step = rng.integers(-3, 4, 32).astype(np.int32)
            x = (x.view(np.int32) + step).view(np.float32)      # a few ulps around the boundary
            x[0] = amax
            X[m, 32 * b: 32 * b + 32] = x
    Wnp = O.random_quant_bytes(O.Q8_0, 4, K, rng)
    _, sums = ops.mul_mat_vec_q(O.Q8_0, torch.from_numpy(Wnp.reshape(-1)).cuda(), torch.from_numpy(X).cuda(), 4, K, dbg=True)
    torch.cuda.synchronize()
    for m in range(8):
        q = O.quantize_row_q8_1_cuda(X[m], K)
        for n in range(4):
            si, _ = O.block_sums(O.Q8_0, Wnp[n], q, K)
            assert np.array_equal(sums[m, n].cpu().numpy(), si)
