"""GPU parity of the round-2 format widening: q4_1 / q5_0 / q5_1 (every path) and q2_K / q3_K / iq4_nl / iq4_xs (decode mat-vec,
dequantisation; reference ggml-quants.c:706-736, 1050-1098, 2428-2467, vecdotq.cuh:218-243, 301-331) through the C ABI against the CPU oracle, to the
bars of test_gpu_mmvq.py / test_gpu_mmq.py: int32 block sums bit-exact (reference ggml-cuda/vecdotq.cuh:49-150), dequantised
rows bit-exact (ggml-quants.c:269-341), fp32 mat-vec within summation-order noise of the oracle's double accumulation, the
tensor-core mat-mul within NMSE 1e-4 of dequantise-then-f32 (the reference's bar for MUL_MAT is 5e-4)."""
import os

import numpy as np
import pytest
import torch

import oracle as O

pytestmark = pytest.mark.gpu
G = dict(np.load(os.path.join(os.path.dirname(__file__), "golden", "quant_golden.npz")))
G.update(np.load(os.path.join(os.path.dirname(__file__), "golden", "quant_golden_legacy.npz")))
NAMES = list(O.LEGACY_TYPES) + list(O.SMALL_K_TYPES)          # tensor-core MMQ (q2_K / q3_K / iq4_xs: shifted 16-byte windows, K % 2048 == 0)
NAMES_MV = NAMES                                               # decode mat-vec, dequantisation, GET_ROWS


def _k_for(name, k):
    return (k + 2047) // 2048 * 2048 if name in ("q2_K", "q3_K", "iq4_xs") else k
TYPES = {**O.LEGACY_TYPES, **O.SMALL_K_TYPES}


def _ops():
    from llamacog_b200 import ops
    return ops


def _run(t, Wnp, Xnp, N, K, **kw):
    ops = _ops()
    W = torch.from_numpy(np.ascontiguousarray(Wnp).reshape(-1)).cuda()
    X = torch.from_numpy(np.ascontiguousarray(Xnp)).cuda()
    out = ops.mul_mat_vec_q(t, W, X, N, K, **kw)
    torch.cuda.synchronize()
    return out


def _close(got, want, tol=2e-5):
    rms = float(np.sqrt((want.astype(np.float64) ** 2).mean())) + 1e-30
    err = float(np.abs(got.astype(np.float64) - want).max()) / rms
    assert err < tol, "max err / rms = %g" % err


def _nmse(got, ref):
    return float(((got.astype(np.float64) - ref) ** 2).sum() / (ref.astype(np.float64) ** 2).sum())


@pytest.mark.parametrize("name", NAMES_MV)
def test_golden_block_sums_bit_exact(name):
    t = TYPES[name]
    dst, sums = _run(t, G["W_" + name], G["X"], 4, 512, dbg=True)
    assert np.array_equal(sums.cpu().numpy(), G["sumi_" + name])
    _close(dst.cpu().numpy(), O.mul_mat_vec_q(t, G["W_" + name], G["X"], 4, 512, 2))


@pytest.mark.parametrize("name", NAMES_MV)
@pytest.mark.parametrize("M", [1, 2, 3, 5, 8])
def test_small_shapes_vs_oracle(name, M):
    t = TYPES[name]
    rng = np.random.default_rng(200 + M)
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


@pytest.mark.parametrize("name", NAMES_MV)
@pytest.mark.parametrize("N,K,M", [(257, 4096, 1), (130, 14336, 1), (96, 4096, 4), (64, 2048, 8), (40, 1024, 7), (33, 8192, 2), (1, 512, 1),
                                   (19, 160, 3), (5, 32, 1)])
def test_model_shapes_vs_oracle(name, N, K, M):
    """K = 160 and 32: rows that end inside a 256-weight item (the tail item holds fewer than 8 blocks)."""
    t = TYPES[name]
    if K % O.BLOCK_ELEMS[t]:
        pytest.skip("K is not a multiple of the super-block")
    rng = np.random.default_rng(N * 7 + K + M)
    W = O.random_quant_bytes(t, N, K, rng)
    X = rng.uniform(-1, 1, (M, K)).astype(np.float32)
    X[0, :32] = 0.0
    dst = _run(t, W, X, N, K).cpu().numpy()
    _close(dst, O.mul_mat_vec_q(t, W, X, N, K, M))
    f32 = O.mul_mat_f32(t, W, X, N, K, M)
    if K >= 512:
        assert _nmse(dst, f32) < 5e-4


@pytest.mark.parametrize("name", NAMES_MV)
def test_strided_unaligned_rows(name):
    ops = _ops()
    t = TYPES[name]
    rng = np.random.default_rng(5)
    N, K, M = 19, 512, 2
    rb = O.row_bytes(t, K)
    nb1 = rb + O.BLOCK_BYTES[t] * 3
    off = O.BLOCK_BYTES[t]
    Wrows = O.random_quant_bytes(t, N, K, rng)
    buf = np.zeros(off + N * nb1 + 64, dtype=np.uint8)
    for n in range(N):
        buf[off + n * nb1: off + n * nb1 + rb] = Wrows[n]
    X = rng.uniform(-1, 1, (M, K)).astype(np.float32)
    Wd = torch.from_numpy(buf).cuda()
    dst = ops.mul_mat_vec_q(t, Wd[off:], torch.from_numpy(X).cuda(), N, K, nb1=nb1)
    torch.cuda.synchronize()
    _close(dst.cpu().numpy(), O.mul_mat_vec_q(t, Wrows, X, N, K, M))


@pytest.mark.parametrize("name", NAMES_MV)
def test_dequantize_bit_exact(name):
    ops = _ops()
    t = TYPES[name]
    for src, ref in (("W_", "deq_"), ("R_", "rdeq_")):
        got = ops.dequantize_rows(t, torch.from_numpy(G[src + name].reshape(-1).copy()).cuda(), 4, 512).cpu().numpy()
        assert np.array_equal(got.view(np.uint32), G[ref + name])
    rng = np.random.default_rng(9)
    raw = O.random_quant_bytes(t, 8, 4096, rng)
    got = ops.dequantize_rows(t, torch.from_numpy(raw.reshape(-1).copy()).cuda(), 8, 4096).cpu().numpy()
    assert np.array_equal(got.view(np.uint32), O.dequantize(t, raw, 8, 4096).view(np.uint32))


@pytest.mark.parametrize("name", NAMES_MV)
def test_prequantised_path_matches_fused(name):
    ops = _ops()
    t = TYPES[name]
    N, K, M = 37, 14336, 2
    rng = np.random.default_rng(K + M)
    W = torch.from_numpy(O.random_quant_bytes(t, N, K, rng).reshape(-1)).cuda()
    X = torch.from_numpy(rng.uniform(-1, 1, (M, K)).astype(np.float32)).cuda()
    d0, s0 = ops.mul_mat_vec_q(t, W, X, N, K, dbg=True)
    d1, s1 = ops.mul_mat_vec_q(t, W, X, N, K, dbg=True, Xq=ops.quantize_acts(X))
    torch.cuda.synchronize()
    assert torch.equal(s0, s1) and torch.equal(d0, d1)


def _dev(Wnp):
    flat = np.ascontiguousarray(Wnp).reshape(-1)
    buf = torch.zeros(flat.size + 256, dtype=torch.uint8, device="cuda")
    buf[:flat.size] = torch.from_numpy(flat).cuda()
    return buf


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("tn", [32, 64, 128, 256])
def test_mmq_one_hot_reads_back_dequantised_weights(name, tn, monkeypatch):
    """One-hot activations read every dequantised weight back through the tensor-core path, element by element."""
    t = TYPES[name]
    N, K = 130, _k_for(name, 512)
    rng = np.random.default_rng(11)
    W = O.random_quant_bytes(t, N, K, rng)
    want = O.dequantize(t, W, N, K)
    monkeypatch.setenv("B200_MMQ_TN", str(tn))
    got = _ops().mul_mat_q(t, _dev(W), torch.eye(K, dtype=torch.float32, device="cuda"), N, K).cpu().numpy().T
    tol = np.abs(want) * 2.0 ** -6 + np.abs(want).max() * 2.0 ** -7
    assert (np.abs(got - want) <= tol).all(), float(np.abs(got - want).max())


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("N,K,M", [(16, 256, 9), (16, 256, 32), (128, 512, 33), (257, 1024, 100), (129, 4096, 129), (512, 2048, 300), (96, 14336, 64)])
def test_mmq_shapes_vs_oracle(name, N, K, M):
    t = TYPES[name]
    K = _k_for(name, K)
    rng = np.random.default_rng(N * 3 + K + M)
    W = O.random_quant_bytes(t, N, K, rng)
    X = rng.uniform(-1, 1, (M, K)).astype(np.float32)
    got = _ops().mul_mat_q(t, _dev(W), torch.from_numpy(X).cuda(), N, K).cpu().numpy()
    ref = O.mul_mat_f32(t, W, X, N, K, M)
    assert _nmse(got, ref) < 1e-4, _nmse(got, ref)


@pytest.mark.parametrize("name", NAMES)
def test_mmq_on_quantiser_output(name):
    """Weights as the reference's quantiser writes them (tests/golden: ggml_quantize_chunk), tiled to a full row tile."""
    t = TYPES[name]
    if _k_for(name, 512) != 512:
        pytest.skip("the golden rows hold two blocks; this format's tensor-core path takes rows of whole 8-block periods")
    N, K, M = 4, 512, 40
    W = G["W_" + name]
    rng = np.random.default_rng(1)
    X = rng.uniform(-1, 1, (M, K)).astype(np.float32)
    got = _ops().mul_mat_q(t, _dev(W), torch.from_numpy(X).cuda(), N, K).cpu().numpy()
    ref = O.mul_mat_f32(t, W, X, N, K, M)
    assert _nmse(got, ref) < 1e-4, _nmse(got, ref)


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("N,K", [(257, 4096), (130, 14336), (33, 8192), (40, 2048), (1500, 4096), (7, 28672)])
def test_register_resident_kernel_vs_oracle(name, N, K):
    """The batch-1 kernel with register-resident activations (mmvq_r.cuh) on the widened formats: every int32 block sum bit-exact
    against the oracle, fp32 within summation-order noise, parity-hook build == production build."""
    ops = _ops()
    t = TYPES[name]
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
    _close(dst.cpu().numpy(), O.mul_mat_vec_q(t, Wnp, Xnp, N, K, 1))
    assert torch.equal(dst, dst2)


@pytest.mark.parametrize("name", NAMES)
def test_register_resident_kernel_is_taken(name):
    """b200_mmvq_fusable says the register-resident kernel takes a 4096 x 4096 launch of every widened format (so the test above and the
    fused decode paths of the plugin really run it, not the general kernel)."""
    import ctypes as C
    ops = _ops()
    t = TYPES[name]
    N = K = 4096
    W = torch.zeros(N * ops.row_bytes(t, K), dtype=torch.uint8, device="cuda")
    d = torch.empty(1, N, device="cuda")
    xq = torch.zeros(K // 256 * 400, dtype=torch.uint8, device="cuda")
    a = ops.MmvqArgs(type=t, W=W.data_ptr(), nb1=ops.row_bytes(t, K), X=None, xs=K, dst=d.data_ptr(), ds=N, N=N, K=K, M=1, ny=1, nz=1, pdl=-1, Xq=xq.data_ptr())
    assert ops.lib().b200_mmvq_fusable(C.byref(a), C.c_int(1), None, None) == 1
