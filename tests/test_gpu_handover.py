"""The row-split hand-over folded into the decode mat-vec (b200_mmvq_args.hs_*, csrc/common.cuh hs_params, csrc/peer.cu),
through the C ABI, on ONE GPU: two streams stand in for the main device and a peer (flags and buffers in the same memory),
so the protocol is pinned on the driver's single-GPU lease too.

  main stream: b200_peer_bcast (activations -> the peer's buffer, raises flag_x) ; own shard with hs_join on the done flag
  peer stream: shard launch(es) with hs_wait = flag_x; the last one lowers flag_x and raises the done flag from its last CTA

Bar: the two slices together equal the unsplit launch bit for bit (rows are independent), every flag and ticket is back at
zero afterwards (so a captured CUDA graph can replay the protocol), the error word stays clear.  Reference counterpart: the
event fork/join + peer memcpys of ggml_cuda_op_mul_mat (ggml-cuda/ggml-cuda.cu:1605-1608, :1628-1630, :1701-1745)."""
import ctypes as C

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


def _launch(ops, t, W, row0, rows, K, xq, dst, N_total, stream, **hs):
    """rows [row0, row0 + rows) of W times the pre-quantised column xq -> dst[row0:row0 + rows]"""
    rb = ops.row_bytes(t, K)
    a = ops.MmvqArgs(type=t, W=W.data_ptr() + row0 * rb, nb1=rb, X=None, xs=K, dst=dst.data_ptr() + 4 * row0, ds=N_total, N=rows, K=K, M=1,
                     ny=1, nz=1, pdl=-1, Xq=xq.data_ptr(), **hs)
    rc = ops.lib().b200_mul_mat_vec_q_ex(C.byref(a), C.c_void_p(stream.cuda_stream))
    assert rc == 0, rc


@pytest.mark.parametrize("tname,N,K,cut", [("q4_K", 4096, 4096, 2048), ("q6_K", 1024, 4096, 512), ("q4_K", 4096, 14336, 1024), ("q5_K", 1024, 8192, 512),
                                           ("q8_0", 512, 4096, 128)])
@pytest.mark.parametrize("repeat", [3])
def test_folded_handover_two_streams(tname, N, K, cut, repeat):
    ops = _ops()
    lib = ops.lib()
    t = O.QUANT_TYPES[tname]
    rng = np.random.default_rng(N + K)
    W = _dev(O.random_quant_bytes(t, N, K, rng))
    main, peer = torch.cuda.Stream(), torch.cuda.Stream()
    flags = torch.zeros(256, dtype=torch.int32, device="cuda")        # [0] flag_x, [32] peer ticket, [64] done flag, [96] main ticket, [128] bcast counter
    err = torch.zeros(32, dtype=torch.int32, device="cuda")
    base = flags.data_ptr()
    flag_x, ticket_p, done, ticket_m, bcast_ctr = base, base + 128, base + 256, base + 384, base + 512
    for it in range(repeat):                                           # several hand-overs in a row: the flags must come back to zero each time
        X = torch.from_numpy(rng.uniform(-1, 1, (1, K)).astype(np.float32)).cuda()
        xq_main = ops.quantize_acts(X)
        xq_peer = torch.zeros_like(xq_main)
        want = ops.mul_mat_vec_q(t, W, X, N, K, Xq=xq_main)
        got = torch.full((1, N), float("nan"), dtype=torch.float32, device="cuda")
        torch.cuda.synchronize()
        dsts = (C.c_void_p * 1)(xq_peer.data_ptr())
        fl = (C.c_void_p * 1)(flag_x)
        assert lib.b200_peer_bcast(C.c_void_p(xq_main.data_ptr()), dsts, fl, 1, C.c_int64(xq_main.numel()), C.c_void_p(bcast_ctr),
                                   C.c_void_p(main.cuda_stream)) == 0
        ev = torch.cuda.Event()
        ev.record(main)
        peer.wait_event(ev)            # one GPU: the spinning CTAs of a waiting launch could keep the broadcast from being scheduled
        # peer: rows [cut, N) in two launches; only the last one carries the last-CTA duties
        mid = cut + (N - cut) // 2 // 2 * 2
        _launch(ops, t, W, cut, mid - cut, K, xq_peer, got, N, peer, hs_wait=flag_x, hs_timeout_ms=5000, hs_error=err.data_ptr())
        _launch(ops, t, W, mid, N - mid, K, xq_peer, got, N, peer, hs_wait=flag_x, hs_clear=flag_x, hs_signal=done, hs_counter=ticket_p,
                hs_timeout_ms=5000, hs_error=err.data_ptr())
        # main: rows [0, cut); its last CTA waits for the peer's done flag
        _launch(ops, t, W, 0, cut, K, xq_main, got, N, main, hs_join=done, hs_join_stride=32, hs_join_mask=1, hs_counter=ticket_m,
                hs_timeout_ms=5000, hs_error=err.data_ptr())
        main.synchronize()             # the end of main's kernel IS the join: the peer's slice must be there without synchronising its stream
        assert torch.equal(got, want), (it, int((got != want).sum()))
        peer.synchronize()
        assert int(flags.abs().sum()) == 0, flags.nonzero().flatten().tolist()
        assert int(err[0]) == 0


def test_handover_timeout_sets_the_error_word_instead_of_trapping():
    """A wait that nobody answers: after the budget the kernel sets the error word and carries on (no trap, no sticky context error)."""
    ops = _ops()
    t = O.QUANT_TYPES["q4_K"]
    N, K = 256, 1024
    rng = np.random.default_rng(5)
    W = _dev(O.random_quant_bytes(t, N, K, rng))
    X = torch.from_numpy(rng.uniform(-1, 1, (1, K)).astype(np.float32)).cuda()
    xq = ops.quantize_acts(X)
    flags = torch.zeros(64, dtype=torch.int32, device="cuda")
    err = torch.zeros(32, dtype=torch.int32, device="cuda")
    got = torch.zeros((1, N), dtype=torch.float32, device="cuda")
    _launch(ops, t, W, 0, N, K, xq, got, N, torch.cuda.current_stream(), hs_wait=flags.data_ptr(), hs_timeout_ms=50, hs_error=err.data_ptr())
    torch.cuda.synchronize()           # would raise on a trap
    assert int(err[0]) == 1
    want = ops.mul_mat_vec_q(t, W, X, N, K, Xq=xq)
    assert torch.equal(got, want)      # the activations were there all along, only the flag was missing
    assert ops.lib().b200_peer_wait_ex(C.c_void_p(flags.data_ptr()), 0, 1, 20, C.c_void_p(err.data_ptr() + 4), None) == 0
    torch.cuda.synchronize()
    assert int(err[1]) == 1
