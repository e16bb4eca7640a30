"""Row split (`--split-mode row`, reference ggml-cuda/ggml-cuda.cu:753-1084, :1428-1746): host-side logic on CPU.

  * the shard geometry the plugin's split buffer type uses (ggml_backend_b200_split_rows) against the rule the
    reference states in get_row_split (ggml-cuda.cu:768-781): boundaries proportional to tensor_split, rounded
    down to 128 rows, the last device takes the remainder;
  * the algorithm itself with two gloo ranks standing in for two devices: every rank takes its row range from the PLUGIN's
    geometry export, multiplies its shard, the slices are gathered, and the result must equal the unsplit product bit for bit
    (rows are independent: no reduction crosses the shard boundary).  The shard product is the product code -- the CUDA mat-vec
    through the C ABI (b200_mul_mat_vec_q) -- in the `gpu` variant (both ranks share cuda:0, so it runs on a one-GPU box too);
    without a GPU the oracle's restatement stands in for the kernel and only the geometry + gather logic is product code.

The reference holds no test for its split buffers ("parity unpinned by the reference"); the GPU side
is pinned by tests/test_gpu_plugin.py::test_row_split_logits_equal_single_gpu on a multi-GPU box."""
import ctypes as C
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PLUGIN = os.path.join(ROOT, "llamacog_b200", "libggml-b200.so")


def _split_rows_ref(n_rows, split):
    """get_row_split of the reference, restated: cumulative fractions, rounded down to 128."""
    n = len(split)
    all_zero = not any(split)
    cum, tot = [], np.float32(0.0)
    for d in range(n):
        cum.append(tot)
        tot = np.float32(tot + np.float32(1.0 if all_zero else split[d]))
    rows = []
    for d in range(n):
        r = int(float(n_rows) * float(np.float32(cum[d] / tot)))
        rows.append(r - r % 128)
    for d in range(1, n):
        rows[d] = max(rows[d], rows[d - 1])
    return rows + [n_rows]


def _plugin():
    if not os.path.exists(PLUGIN):
        pytest.skip("plugin not built (needs the reference headers)")
    C.CDLL(os.path.join(ROOT, "llamacog_b200", "libb200ops.so"), mode=C.RTLD_GLOBAL)
    C.CDLL(os.path.join(ROOT, "baseline", "_ref", "lib", "libggml-base.so"), mode=C.RTLD_GLOBAL)
    return C.CDLL(PLUGIN)


def _split_rows(lib, n_rows, split):
    n = len(split)
    ts = (C.c_float * n)(*split)
    out = (C.c_longlong * (n + 1))()
    lib.ggml_backend_b200_split_rows(C.c_longlong(n_rows), C.c_int(n), ts, out)
    return list(out)


@pytest.mark.parametrize("n_rows", [1, 127, 128, 1024, 4096, 8192, 14336, 28672, 128256])
@pytest.mark.parametrize("split", [[0, 0], [1, 1], [3, 1], [1, 0], [0, 1], [1, 1, 1, 1], [0, 0, 0, 0, 0, 0, 0, 0], [5, 1, 1, 1, 0, 0, 0, 0]])
def test_split_rows_geometry(n_rows, split):
    lib = _plugin()
    got = _split_rows(lib, n_rows, split)
    n = len(split)
    assert got[0] == 0 and got[n] == n_rows
    assert all(got[d] <= got[d + 1] for d in range(n))                # shards tile the matrix without overlap
    assert all(got[d] % 128 == 0 for d in range(n))                   # interior boundaries on MMQ tiles
    assert got == _split_rows_ref(n_rows, [float(s) for s in split])
    for d in range(n):                                                 # a zero share holds no rows (unless it is the last device's remainder)
        if any(split) and split[d] == 0 and d < n - 1:
            assert got[d + 1] == got[d]


def _rank_main(rank, world, port, t, N, K, M, split, ret, use_gpu=False):
    import torch.distributed as dist
    import oracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import torch
    rng = np.random.default_rng(7)                                    # same seed on every rank: same W and X
    W = O.random_quant_bytes(t, N, K, rng)
    X = rng.standard_normal((M, K)).astype(np.float32)
    rows = _split_rows(_plugin(), N, split) if os.path.exists(PLUGIN) else _split_rows_ref(N, split)      # the plugin's own geometry
    r0, r1 = rows[rank], rows[rank + 1]
    rb = O.row_bytes(t, K)
    Wb = np.ascontiguousarray(W).reshape(N, rb)

    def product(Wrows, n):
        if n == 0:
            return np.zeros((M, 0), np.float32)
        if use_gpu:                                                   # product code: the CUDA mat-vec through the C ABI
            from llamacog_b200 import ops
            Wd = torch.zeros(Wrows.size + 256, dtype=torch.uint8, device="cuda")
            Wd[:Wrows.size] = torch.from_numpy(np.ascontiguousarray(Wrows).reshape(-1)).cuda()
            out = ops.mul_mat_vec_q(t, Wd, torch.from_numpy(X).cuda(), n, K)
            torch.cuda.synchronize()
            return out.cpu().numpy()
        return O.mul_mat_vec_q(t, Wrows.copy(), X, n, K, M)
    mine = product(Wb[r0:r1], r1 - r0)
    # the gather: every rank contributes its [M, rows] slice (padded to the widest shard for all_gather)
    width = max(rows[d + 1] - rows[d] for d in range(world))
    pad = np.zeros((M, width), np.float32)
    pad[:, :r1 - r0] = mine
    parts = [torch.zeros(M, width) for _ in range(world)]
    dist.all_gather(parts, torch.from_numpy(pad))
    full = np.concatenate([parts[d].numpy()[:, :rows[d + 1] - rows[d]] for d in range(world)], axis=1)
    if rank == 0:
        want = product(Wb, N)                                         # the unsplit product by the same code
        ok = bool(np.array_equal(full.view(np.uint32), want.view(np.uint32)))
        if use_gpu:                                                   # ... which itself is held to the oracle
            ref = O.mul_mat_vec_q(t, Wb, X, N, K, M).astype(np.float64)
            ok = ok and float(np.abs(want - ref).max()) <= 2e-5 * float(np.sqrt((ref ** 2).mean()) + 1e-30) * 4
        ret.put((ok, full.shape))
    dist.barrier()
    dist.destroy_process_group()


CASES = [("q4_K", 384, 512, 1, [1.0, 1.0]), ("q6_K", 640, 256, 3, [3.0, 1.0]), ("q4_0", 256, 256, 2, [1.0, 0.0])]


@pytest.mark.gpu
@pytest.mark.parametrize("tname,N,K,M,split", CASES + [("q4_K", 4096, 4096, 1, [1.0, 1.0]), ("q5_K", 1024, 8192, 1, [1.0, 1.0])])
def test_row_sharded_matmul_two_ranks_gloo_cuda_kernels(tname, N, K, M, split):
    """The same with the CUDA mat-vec of the product as every rank's shard multiply (C ABI, both ranks on cuda:0)."""
    test_row_sharded_matmul_two_ranks_gloo(tname, N, K, M, split, use_gpu=True)


@pytest.mark.parametrize("tname,N,K,M,split", CASES)
def test_row_sharded_matmul_two_ranks_gloo(tname, N, K, M, split, use_gpu=False):
    import torch.multiprocessing as mp
    import oracle as O
    t = O.TYPES[tname] if hasattr(O, "TYPES") else {"q4_0": 2, "q8_0": 8, "q4_K": 12, "q5_K": 13, "q6_K": 14}[tname]
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, t, N, K, M, split, ret, use_gpu)) for r in range(2)]
    for p in procs:
        p.start()
    ok, shape = ret.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert shape == (M, N)
    assert ok, "gathered shards differ from the unsplit product"
