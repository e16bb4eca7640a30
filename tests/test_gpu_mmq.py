"""GPU parity of the tensor-core prefill mat-mul (b200_mul_mat_q, through the C ABI) against the CPU oracle.

Bar: NMSE <= 5e-4 against dequantise-then-f32 matmul, the reference's own bar for MUL_MAT
(tests/test-backend-ops.cpp:1990-1992); the tests hold the kernel to 1e-4.  One-hot activations read the
dequantised weights back through the MMA and are checked element-wise against dequantize_row_* (oracle)."""
import os

import numpy as np
import pytest
import torch

import oracle as O

pytestmark = pytest.mark.gpu
NAMES = list(O.QUANT_TYPES)
NMSE_BAR = 1e-4


def _ops():
    from llamacog_b200 import ops
    return ops


def _dev(Wnp):
    flat = np.ascontiguousarray(Wnp).reshape(-1)
    buf = torch.zeros(flat.size + 256, dtype=torch.uint8, device="cuda")      # slack: the Q6_K box reads 30 bytes past a block
    buf[:flat.size] = torch.from_numpy(flat).cuda()
    return buf


def _k_for(name, k):
    return (k + 2047) // 2048 * 2048 if name == "q6_K" else k       # Q6_K rows must be a multiple of 16 bytes for TMA


def _nmse(got, ref):
    return float(((got.astype(np.float64) - ref) ** 2).sum() / (ref.astype(np.float64) ** 2).sum())


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("tn", [32, 64, 128, 256])
def test_one_hot_reads_back_dequantised_weights(name, tn, monkeypatch):
    t = O.QUANT_TYPES[name]
    N, K = 130, _k_for(name, 512)
    rng = np.random.default_rng(11)
    W = O.random_quant_bytes(t, N, K, rng)
    want = O.dequantize(t, W, N, K)
    monkeypatch.setenv("B200_MMQ_TN", str(tn))
    got = _ops().mul_mat_q(t, _dev(W), torch.eye(K, dtype=torch.float32, device="cuda"), N, K).cpu().numpy().T
    tol = np.abs(want) * 2.0 ** -6 + np.abs(want).max() * 2.0 ** -7
    assert (np.abs(got - want) <= tol).all(), float(np.abs(got - want).max())


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("N,K,M", [(16, 256, 9), (16, 256, 16), (16, 256, 32), (128, 512, 33), (257, 1024, 100), (129, 4096, 129),
                                   (512, 2048, 300), (96, 14336, 64)])
def test_shapes_vs_oracle(name, N, K, M):
    """n = 9..32 at m=16,k=256 is the reference's own grid (tests/test-backend-ops.cpp:4246-4266); the rest are ragged tiles."""
    t = O.QUANT_TYPES[name]
    K = _k_for(name, K)
    rng = np.random.default_rng(N * 3 + K + M)
    W = O.random_quant_bytes(t, N, K, rng)
    X = rng.uniform(-1, 1, (M, K)).astype(np.float32)
    got = _ops().mul_mat_q(t, _dev(W), torch.from_numpy(X).cuda(), N, K).cpu().numpy()
    ref = O.mul_mat_f32(t, W, X, N, K, M)
    assert _nmse(got, ref) < NMSE_BAR, _nmse(got, ref)


@pytest.mark.parametrize("name", NAMES)
def test_strided_activations_and_padded_rows(name):
    """src1 rows with a stride (a view) and weight rows with a larger nb1."""
    t = O.QUANT_TYPES[name]
    N, K, M = 70, _k_for(name, 768), 40
    rng = np.random.default_rng(3)
    W = O.random_quant_bytes(t, N, K, rng)
    rb = W.shape[1]
    nb1 = rb + 64
    Wp = np.zeros((N, nb1), dtype=np.uint8); Wp[:, :rb] = W
    Xfull = rng.uniform(-1, 1, (M, K + 12)).astype(np.float32)
    Xd = torch.from_numpy(Xfull).cuda()[:, 4:4 + K]                    # 16-byte aligned start, row stride K+12
    got = _ops().mul_mat_q(t, _dev(Wp), Xd, N, K, nb1=nb1).cpu().numpy()
    ref = O.mul_mat_f32(t, W, np.ascontiguousarray(Xfull[:, 4:4 + K]), N, K, M)
    assert _nmse(got, ref) < NMSE_BAR
    Xu = torch.from_numpy(Xfull).cuda()[:, 1:1 + K]                    # unaligned start: scalar conversion path
    got = _ops().mul_mat_q(t, _dev(Wp), Xu, N, K, nb1=nb1).cpu().numpy()
    ref = O.mul_mat_f32(t, W, np.ascontiguousarray(Xfull[:, 1:1 + K]), N, K, M)
    assert _nmse(got, ref) < NMSE_BAR


@pytest.mark.parametrize("name", ["q4_K", "q6_K"])
def test_model_size_properties(name):
    """Llama-3-8B FFN shape at pp512: agreement with the GEMV path (itself oracle-checked), exact scaling by a
    power of two, and determinism."""
    ops = _ops()
    t = O.QUANT_TYPES[name]
    N, K, M = 14336, 4096, 512
    rng = np.random.default_rng(9)
    one = O.random_quant_bytes(t, 256, K, rng)
    W = _dev(np.tile(one, (N // 256, 1)))
    X = torch.randn(M, K, device="cuda")
    a = ops.mul_mat_q(t, W, X, N, K)
    b = ops.mul_mat_q(t, W, X, N, K)
    assert torch.equal(a, b)
    assert torch.equal(ops.mul_mat_q(t, W, X * 4.0, N, K), a * 4.0)
    assert torch.equal(a[:, :256], a[:, 256:512])                       # tiled weights -> tiled outputs
    gemv = torch.cat([ops.mul_mat_vec_q(t, W, X[i:i + 8].contiguous(), N, K) for i in range(0, 64, 8)])
    num = ((a[:64] - gemv) ** 2).sum().item(); den = (gemv ** 2).sum().item()
    assert num / den < 2e-4, num / den


def test_unsupported_shapes_are_refused():
    ops = _ops()
    W = torch.zeros(4096, dtype=torch.uint8, device="cuda")
    X = torch.zeros(16, 128, device="cuda")
    with pytest.raises(RuntimeError):
        ops.mul_mat_q(O.QUANT_TYPES["q4_0"], W, X, 4, 128)              # K % 256 != 0
    with pytest.raises(RuntimeError):
        ops.mul_mat_q(O.QUANT_TYPES["q6_K"], W, torch.zeros(16, 256, device="cuda"), 4, 256)      # 210-byte row stride


@pytest.mark.parametrize("name", ["q4_K", "q6_K", "q8_0", "q4_0", "q5_K"])
@pytest.mark.parametrize("N,K,M", [(4096, 4096, 9), (1024, 14336, 32), (300, 8192, 17), (2048, 2048, 24)])
def test_k_split_single_token_tile(name, N, K, M):
    """Launches of one 32-token tile (batched decode steps, short prompts) whose grid would leave most SMs idle are split along K
    (mmq.cu: plan_ksplit, slices summed in fixed order by ksplit_reduce_kernel): 2 .. 8 slices here, Q6_K in whole 8-block periods."""
    t = O.QUANT_TYPES[name]
    K = _k_for(name, K)
    rng = np.random.default_rng(N + K + M)
    W = O.random_quant_bytes(t, N, K, rng)
    X = rng.uniform(-1, 1, (M, K)).astype(np.float32)
    ops = _ops()
    got = ops.mul_mat_q(t, _dev(W), torch.from_numpy(X).cuda(), N, K).cpu().numpy()
    again = ops.mul_mat_q(t, _dev(W), torch.from_numpy(X).cuda(), N, K).cpu().numpy()
    assert np.array_equal(got.view(np.uint32), again.view(np.uint32))      # deterministic
    rows = rng.choice(N, min(N, 96), replace=False)                        # the oracle on a sample of the rows
    ref = O.mul_mat_f32(t, np.ascontiguousarray(W.reshape(N, -1)[rows]), X, len(rows), K, M)
    assert _nmse(got[:, rows], ref) < NMSE_BAR, _nmse(got[:, rows], ref)


@pytest.mark.parametrize("name", ["q4_K", "q6_K"])
def test_k_split_group_launch(name):
    """q/k/v of a 12-token step in one split grid: every matrix against the oracle."""
    ops = _ops()
    t = O.QUANT_TYPES[name]
    K, M, Ns = 4096, 12, (4096, 1024, 512)
    rng = np.random.default_rng(5)
    Wn = [O.random_quant_bytes(t, n, K, rng) for n in Ns]
    X = rng.uniform(-1, 1, (M, K)).astype(np.float32)
    Wd = [_dev(w) for w in Wn]
    Xd = torch.from_numpy(X).cuda()
    got = ops.mul_mat_q_group(t, Wd, list(Ns), Xd, K)
    for g, n in enumerate(Ns):
        rows = np.arange(0, n, 7)
        ref = O.mul_mat_f32(t, np.ascontiguousarray(Wn[g].reshape(n, -1)[rows]), X, len(rows), K, M)
        assert _nmse(got[g].cpu().numpy()[:, rows], ref) < NMSE_BAR, g
        # every matrix keeps the slice count of its own launch: grouped == single, bit for bit (fused and unfused graphs agree)
        assert torch.equal(got[g], ops.mul_mat_q(t, Wd[g], Xd, n, K)), g


@pytest.mark.parametrize("name", ["q4_K", "q6_K", "q8_0"])
@pytest.mark.parametrize("Ns,M", [((4096, 1024, 1024), 512), ((512, 200), 77), ((1024, 1024, 384), 300)])
def test_group_launch_equals_single_launches(name, Ns, M):
    """q/k/v of a prefill batch in one grid (b200_mul_mat_q_bf16_group): each matrix's result must equal, bit for bit, its own
    launch (same K-order of the f32 accumulation whatever the token tile), ragged last row tiles included."""
    ops = _ops()
    t = O.QUANT_TYPES[name]
    K = 2048
    rng = np.random.default_rng(len(Ns) * 1000 + M)
    Ws = [_dev(O.random_quant_bytes(t, n, K, rng)) for n in Ns]
    torch.manual_seed(M)
    X = torch.randn(M, K, device="cuda")
    got = ops.mul_mat_q_group(t, Ws, list(Ns), X, K)
    for g, n in enumerate(Ns):
        want = ops.mul_mat_q(t, Ws[g], X, n, K)
        assert torch.equal(got[g], want), "matrix %d of the group differs" % g
