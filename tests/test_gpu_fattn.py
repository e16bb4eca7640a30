"""GPU parity of FLASH_ATTN_EXT (through the C ABI) against a plain PyTorch fp32 reference of the same op
(softmax(scale * Q K^T + mask) V, reference ggml/include/ggml.h:1848-1870, ggml-cpu/ops.cpp:7000-7250).

Tolerance: NMSE <= 5e-4 is the reference's bar for this op (tests/test-backend-ops.cpp:3340-3347); the tests hold
both the vector kernel (<= 8 rows) and the tensor-core kernel (> 8 rows) to 1e-5 with f16 KV."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ops():
    from llamacog_b200 import ops
    return ops


def _ref(q, k, v, mask, scale, softcap=0.0):
    ne3, n_head, N, D = q.shape
    rep = n_head // k.shape[1]
    kf = k.float().repeat_interleave(rep, dim=1); vf = v.float().repeat_interleave(rep, dim=1)
    s = torch.einsum("bhqd,bhkd->bhqk", q.half().float(), kf) * scale
    if softcap:
        s = softcap * torch.tanh(s / softcap)
    if mask is not None:
        s = s + mask[:N, :k.shape[2]].float()[None, None]
    p = torch.softmax(s, dim=-1)
    return torch.einsum("bhqk,bhkd->bqhd", p, vf)


def _nmse(a, b):
    return float(((a.double() - b.double()) ** 2).sum() / (b.double() ** 2).sum())


def _causal_mask(N, n_kv, n_past, pad_q=64, pad_kv=256):
    Np = (N + pad_q - 1) // pad_q * pad_q; Kp = (n_kv + pad_kv - 1) // pad_kv * pad_kv
    m = torch.full((Np, Kp), float("-inf"), dtype=torch.float16, device="cuda")
    for i in range(N):
        m[i, : n_past + i + 1] = 0
    return m


@pytest.mark.parametrize("D", [64, 128])
@pytest.mark.parametrize("N,n_kv,n_past", [(1, 300, 299), (8, 512, 504), (9, 256, 247), (32, 512, 480), (35, 1024, 100), (128, 128, 0),
                                            (200, 200, 0), (512, 512, 0), (130, 700, 570), (512, 2048, 1536), (1, 4100, 4099), (3, 2000, 1990),
                                            (5, 1500, 700)])
def test_causal_gqa(D, N, n_kv, n_past):
    torch.manual_seed(N + n_kv + D)
    n_head, n_head_kv = 8, 2
    q = torch.randn(1, n_head, N, D, device="cuda")
    k = torch.randn(1, n_head_kv, n_kv, D, device="cuda").half()
    v = torch.randn(1, n_head_kv, n_kv, D, device="cuda").half()
    mask = _causal_mask(N, n_kv, n_past)
    scale = 1.0 / math.sqrt(D)
    out = _ops().flash_attn_ext(q, k, v, mask, scale)
    assert _nmse(out, _ref(q, k, v, mask, scale)) < 1e-5


@pytest.mark.parametrize("cache", ["f16", "q8_0"])
@pytest.mark.parametrize("N,n_kv,n_head,n_head_kv", [(1, 32768, 32, 8), (1, 32768, 8, 1), (3, 32768, 8, 4), (1, 28672 + 129, 32, 8), (8, 16384, 8, 2)])
def test_decode_at_32k_context(cache, N, n_kv, n_head, n_head_kv):
    """Config 3 of BASELINE.json (32k context): the grouped split-KV decode kernel (K tiles staged through cp.async for an f16 cache,
    up to 128 KV splits + combine) on the shapes of Llama-3-8B and on groups of 8 / 2 queries per KV head, f16 and q8_0 cache,
    against the fp32 reference on the values the cache really holds."""
    import numpy as np
    import oracle as O
    D = 128
    torch.manual_seed(n_kv + N)
    q = torch.randn(1, n_head, N, D, device="cuda")
    kf = torch.randn(1, n_head_kv, n_kv, D); vf = torch.randn(1, n_head_kv, n_kv, D)
    mask = torch.full(((N + 63) // 64 * 64, (n_kv + 255) // 256 * 256), float("-inf"), dtype=torch.float16, device="cuda")
    for i in range(N):
        mask[i, : n_kv - N + i + 1] = 0
    scale = 1.0 / math.sqrt(D)
    if cache == "f16":
        k = kf.half().cuda(); v = vf.half().cuda()
        out = _ops().flash_attn_ext(q, k, v, mask, scale)
        ref = _ref(q, k, v, mask, scale)
        bar = 1e-5
    else:
        t = O.QUANT_TYPES["q8_0"]
        kq = O.quantize_row_q8_0_ref(kf.numpy().reshape(-1)); vq = O.quantize_row_q8_0_ref(vf.numpy().reshape(-1))
        kd = torch.from_numpy(O.dequantize(t, kq, n_head_kv * n_kv, D)).reshape(1, n_head_kv, n_kv, D).cuda()
        vd = torch.from_numpy(O.dequantize(t, vq, n_head_kv * n_kv, D)).reshape(1, n_head_kv, n_kv, D).cuda()
        k8 = torch.from_numpy(kq).reshape(1, n_head_kv, n_kv, D // 32 * 34).cuda(); v8 = torch.from_numpy(vq).reshape(1, n_head_kv, n_kv, D // 32 * 34).cuda()
        out = _ops().flash_attn_ext(q, k8, v8, mask, scale)
        ref = _ref(q, kd, vd, mask, scale)
        bar = 5e-4                      # Q is quantised to int8 for the q8_0 K dot (as the reference does): the op's own bar
    assert _nmse(out, ref) < bar, _nmse(out, ref)


@pytest.mark.parametrize("N", [4, 48, 256])
def test_no_mask_softcap_and_batch(N):
    torch.manual_seed(N)
    D, n_head, n_kv = 128, 4, 333
    q = torch.randn(2, n_head, N, D, device="cuda")
    k = torch.randn(2, n_head, n_kv, D, device="cuda").half()
    v = torch.randn(2, n_head, n_kv, D, device="cuda").half()
    out = _ops().flash_attn_ext(q, k, v, None, 0.11)
    assert _nmse(out, _ref(q, k, v, None, 0.11)) < 1e-5
    out = _ops().flash_attn_ext(q, k, v, None, 0.11, logit_softcap=10.0)
    assert _nmse(out, _ref(q, k, v, None, 0.11, softcap=10.0)) < 1e-5


def test_kv_cache_view_layout():
    """K and V as llama.cpp hands them over: [n_kv][n_head_kv][D] memory viewed as [n_head_kv][n_kv][D] (head stride < row stride),
    random mask with whole 128-blocks masked out (other sequences), and rows that attend to nothing in some tiles."""
    torch.manual_seed(5)
    D, n_head, n_head_kv, N, n_kv = 128, 8, 2, 256, 1024
    q = torch.randn(1, n_head, N, D, device="cuda")
    kc = torch.randn(1, n_kv, n_head_kv, D, device="cuda").half(); vc = torch.randn(1, n_kv, n_head_kv, D, device="cuda").half()
    k = kc.permute(0, 2, 1, 3); v = vc.permute(0, 2, 1, 3)
    mask = torch.zeros(N, n_kv, dtype=torch.float16, device="cuda")
    mask[:, 256:512] = float("-inf")
    mask[:128, 640:768] = float("-inf")
    mask[torch.rand(N, n_kv, device="cuda") < 0.3] = float("-inf")
    mask[:, 0] = 0
    out = _ops().flash_attn_ext(q, k, v, mask, 0.09)
    assert _nmse(out, _ref(q, k, v, mask, 0.09)) < 1e-5


@pytest.mark.parametrize("D", [64, 128])
@pytest.mark.parametrize("N,n_kv,n_past", [(1, 300, 299), (8, 512, 504), (9, 256, 247), (64, 512, 448), (130, 700, 570), (512, 2048, 1536),
                                            (1, 4100, 4099), (3, 2000, 1990)])
def test_q8_0_kv_cache(D, N, n_kv, n_past):
    """K and V stored as q8_0 (-ctk q8_0 -ctv q8_0): decode rows take the vector kernel (q8_1 x q8_0 integer dot products), 9 rows and
    more dequantise the cache to f16 and take the tensor-core kernel.  Reference: fattn-vec-f32.cuh:72-126, fattn-common.cuh:700-760;
    the oracle quantises with quantize_row_q8_0_ref (ggml-quants.c:198-231) and the expectation attends over the dequantised values."""
    import numpy as np
    import oracle as O
    torch.manual_seed(N + n_kv + D + 1)
    n_head, n_head_kv = 8, 2
    q = torch.randn(1, n_head, N, D, device="cuda")
    kf = torch.randn(1, n_head_kv, n_kv, D)
    vf = torch.randn(1, n_head_kv, n_kv, D)
    kq = O.quantize_row_q8_0_ref(kf.numpy().reshape(-1)); vq = O.quantize_row_q8_0_ref(vf.numpy().reshape(-1))
    kd = torch.from_numpy(O.dequantize(O.QUANT_TYPES["q8_0"], kq, n_head_kv * n_kv, D)).reshape(1, n_head_kv, n_kv, D).cuda()
    vd = torch.from_numpy(O.dequantize(O.QUANT_TYPES["q8_0"], vq, n_head_kv * n_kv, D)).reshape(1, n_head_kv, n_kv, D).cuda()
    k8 = torch.from_numpy(kq).reshape(1, n_head_kv, n_kv, D // 32 * 34).cuda()
    v8 = torch.from_numpy(vq).reshape(1, n_head_kv, n_kv, D // 32 * 34).cuda()
    mask = _causal_mask(N, n_kv, n_past)
    scale = 1.0 / math.sqrt(D)
    out = _ops().flash_attn_ext(q, k8, v8, mask, scale)
    ref = _ref(q, kd, vd, mask, scale)
    assert _nmse(out, ref) < 5e-4, _nmse(out, ref)


@pytest.mark.parametrize("D,N,n_kv,n_past", [(128, 512, 512, 0), (64, 200, 200, 0), (128, 130, 700, 570), (128, 35, 1024, 100)])
def test_prefill_bf16_rows_for_the_output_projection(D, N, n_kv, n_past):
    """b200_op_flash_attn_ext_xb: the tensor-core kernel also stores its result as the bf16 activation rows of the MMQ that reads
    it; dst must equal the plain call bit for bit and the rows must be the round-to-nearest bf16 of dst (what
    b200_convert_acts_bf16 would have produced)."""
    import ctypes as C
    ops = _ops(); L = ops.lib()
    torch.manual_seed(N + n_kv + D)
    n_head, n_head_kv = 8, 2
    q = torch.randn(1, n_head, N, D, device="cuda")
    k = torch.randn(1, n_head_kv, n_kv, D, device="cuda").half()
    v = torch.randn(1, n_head_kv, n_kv, D, device="cuda").half()
    mask = _causal_mask(N, n_kv, n_past)
    scale = 1.0 / math.sqrt(D)
    want = ops.flash_attn_ext(q, k, v, mask, scale)
    dst = torch.zeros_like(want)
    xb = torch.zeros(N, n_head * D, dtype=torch.bfloat16, device="cuda")
    tq, tk, tv, tm, td = ops.as_b200(q), ops.as_b200(k), ops.as_b200(v), ops.as_b200(mask), ops.as_b200(dst)
    assert L.b200_flash_attn_xb_supported(C.byref(tq), C.byref(tk), C.byref(td)) == 1
    need = L.b200_flash_attn_scratch_bytes(C.byref(tq), C.byref(tk))
    scratch = torch.empty(max(int(need), 16), dtype=torch.uint8, device="cuda")
    rc = L.b200_op_flash_attn_ext_xb(C.byref(tq), C.byref(tk), C.byref(tv), C.byref(tm), C.byref(td), C.c_float(scale), C.c_float(0.0), C.c_float(0.0),
                                     C.c_void_p(scratch.data_ptr()), C.c_int64(scratch.numel()), C.c_void_p(xb.data_ptr()),
                                     C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.equal(dst, want)
    assert torch.equal(xb.view(torch.int16), dst.view(N, n_head * D).to(torch.bfloat16).view(torch.int16))
    # decode shapes are not this entry point's
    q1 = torch.randn(1, n_head, 1, D, device="cuda"); d1 = torch.zeros(1, 1, n_head, D, device="cuda")
    assert L.b200_flash_attn_xb_supported(C.byref(ops.as_b200(q1)), C.byref(tk), C.byref(ops.as_b200(d1))) == 0
