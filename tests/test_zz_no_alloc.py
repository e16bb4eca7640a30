"""b200_mul_mat_q_noalloc (the allocation-free variant of the tensor-core prefill mat-mul, scratch in a static __device__
array) through the C ABI against the existing entry point b200_mul_mat_q, on two calls in a row.

Collected last on purpose.  Tolerance: the two entry points launch the same kernels on the
same values, so the outputs must be bit-identical -- tighter than the NMSE <= 1e-4 bar tests/test_gpu_mmq.py holds
b200_mul_mat_q to against the oracle, which is checked here as well."""
import numpy as np
import pytest
import torch

import oracle as O

pytestmark = pytest.mark.gpu


def _ops():
    from llamacog_b200 import ops
    return ops


def _dev(Wnp):
    flat = np.ascontiguousarray(Wnp).reshape(-1)
    buf = torch.zeros(flat.size + 256, dtype=torch.uint8, device="cuda")
    buf[:flat.size] = torch.from_numpy(flat).cuda()
    return buf


def _nmse(got, ref):
    return float(((got.astype(np.float64) - ref) ** 2).sum() / (ref.astype(np.float64) ** 2).sum())


def test_noalloc_matches_mul_mat_q_on_two_calls_in_a_row():
    ops = _ops()
    t = O.QUANT_TYPES["q4_K"]
    rng = np.random.default_rng(2024)
    # second call: fewer tokens and a different K, so it runs over the first call's leftovers in the shared array
    for N, K, M in [(257, 1024, 100), (130, 512, 33)]:
        W = O.random_quant_bytes(t, N, K, rng)
        X = rng.uniform(-1, 1, (M, K)).astype(np.float32)
        Wd, Xd = _dev(W), torch.from_numpy(X).cuda()
        want = ops.mul_mat_q(t, Wd, Xd, N, K)
        dst = torch.full((M, N), float("nan"), dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        before = torch.cuda.memory_stats()["allocation.all.allocated"]      # running count of torch's device allocations
        rc, got = ops.mul_mat_q_noalloc(t, Wd, Xd, N, K, dst=dst)
        assert rc == 0, rc
        assert torch.cuda.memory_stats()["allocation.all.allocated"] == before      # the binding did not allocate either
        torch.cuda.synchronize()
        assert torch.equal(got, want), float((got - want).abs().max())
        ref = O.mul_mat_f32(t, W, X, N, K, M)
        assert _nmse(got.cpu().numpy(), ref) < 1e-4


def test_noalloc_rejects_sizes_beyond_the_static_array():
    ops = _ops()
    t = O.QUANT_TYPES["q4_K"]
    N, K = 128, 4096
    M = (8 << 20) // (K * 2) + 1                                  # one token row more than B200_NOALLOC_SCRATCH_BYTES holds
    W = torch.zeros(N * ops.row_bytes(t, K) + 256, dtype=torch.uint8, device="cuda")
    X = torch.zeros((M, K), dtype=torch.float32, device="cuda")
    rc, _ = ops.mul_mat_q_noalloc(t, W, X, N, K)
    assert rc == -7                                               # B200_ERR_NOALLOC_TOO_LARGE
    rc, got = ops.mul_mat_q_noalloc(t, W, X[:M - 1], N, K)        # exactly the maximum is accepted
    assert rc == 0
    torch.cuda.synchronize()
    assert float(got.abs().max()) == 0.0
