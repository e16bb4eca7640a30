#!/usr/bin/env python
"""Random-init GGUF models of the BASELINE.json architectures, quantised with the reference's own
quantiser.

Weights are N(0, 0.02) (fixed seed), quantised by `ggml_quantize_chunk` of the UNMODIFIED reference
libggml-base.so (baseline/_ref/lib; the function llama-quantize calls, ggml/src/ggml.c) with the
tensor-type mix llama-quantize applies for the requested ftype (reference src/llama-quant.cpp:185-187
use_more_bits, :207-224 output/token_embd, :291-353 attn_v / ffn_down, :294-311 70B and 8-expert
rules).  To keep generation to seconds on the GPU box, one layer's tensors are generated and
quantised once and their bytes are written for every layer (each layer is still a distinct region
of the file and of device memory, so a decode step streams the full model from HBM).
Tokenizer: `no_vocab` with llama.vocab_size set (reference src/llama-vocab.cpp:1383-1399); llama-bench
only draws random token ids (tools/llama-bench/llama-bench.cpp:1760-1792); the llama-bpe vocab blob
BASELINE.json names is absent from the reference mount (.MISSING_LARGE_BLOBS).

usage: make_gguf.py <preset> <ftype> <out.gguf>     presets: llama3-8b llama3.2-1b llama3-70b mixtral-8x7b tiny tiny-moe
                                                     ftypes:  q4_k_m q4_0 q4_1 q5_0 q5_1 q8_0 q2_k q3_k iq4_nl iq4_xs f16   (q2_k ... iq4_xs: every 2-D tensor of that type, output Q6_K)
"""
import concurrent.futures as cf
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

PRESETS = {
    #               n_embd n_layer n_head n_kv  n_ff  n_vocab  experts used tied  ctx
    "llama3-8b":    (4096, 32, 32, 8, 14336, 128256, 0, 0, False, 8192),
    "llama3.2-1b":  (2048, 16, 32, 8, 8192, 128256, 0, 0, True, 8192),
    "llama3-70b":   (8192, 80, 64, 8, 28672, 128256, 0, 0, False, 8192),
    "mixtral-8x7b": (4096, 32, 32, 8, 14336, 32000, 8, 2, False, 32768),
    "tiny":         (256, 2, 4, 2, 512, 1024, 0, 0, False, 2048),
    "tiny-moe":     (256, 2, 4, 2, 512, 1024, 4, 2, False, 2048),
}
F32, F16, Q4_0, Q8_0, Q4_K, Q5_K, Q6_K = 0, 1, 2, 8, 12, 13, 14
Q4_1, Q5_0, Q5_1 = 3, 6, 7
Q2_K, Q3_K = 10, 11
IQ4_NL, IQ4_XS = 20, 23
FTYPE_ID = {"f16": 1, "q4_0": 2, "q4_1": 3, "q8_0": 7, "q5_0": 8, "q5_1": 9, "q2_k": 10, "q3_k": 11, "q4_k_m": 15, "iq4_nl": 25, "iq4_xs": 30}      # llama_ftype (include/llama.h)


def use_more_bits(i, n):
    return i < n // 8 or i >= 7 * n // 8 or (i - n // 8) % 3 == 2


def tensor_type(name, il, n_layer, ftype, n_expert, is_70b, tied):
    if name.endswith("_norm.weight") or name == "ffn_gate_inp.weight":
        return F32
    if ftype == "f16":
        return F16
    base = {"q4_0": Q4_0, "q4_1": Q4_1, "q5_0": Q5_0, "q5_1": Q5_1, "q8_0": Q8_0, "q2_k": Q2_K, "q3_k": Q3_K, "iq4_nl": IQ4_NL, "iq4_xs": IQ4_XS, "q4_k_m": Q4_K}[ftype]
    if name == "output.weight" or (tied and name == "token_embd.weight"):
        return Q8_0 if ftype == "q8_0" else Q6_K
    if ftype != "q4_k_m":
        return base
    if name == "attn_v.weight":
        t = Q6_K if use_more_bits(il, n_layer) else Q4_K
        if is_70b and t == Q4_K:
            t = Q5_K
        if n_expert == 8:
            t = Q8_0
        return t
    if name == "attn_k.weight" and n_expert == 8:
        return Q8_0
    if name in ("ffn_down.weight", "ffn_down_exps.weight"):
        return Q6_K if use_more_bits(il, n_layer) else Q4_K
    return base


class Quantizer:
    def __init__(self):
        self.lib = C.CDLL(os.path.join(ROOT, "baseline", "_ref", "lib", "libggml-base.so"))
        self.lib.ggml_quantize_chunk.restype = C.c_size_t
        self.lib.ggml_quantize_chunk.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int64, C.c_void_p]
        self.lib.ggml_row_size.restype = C.c_size_t
        self.lib.ggml_row_size.argtypes = [C.c_int, C.c_int64]
        self.pool = cf.ThreadPoolExecutor(max_workers=os.cpu_count() or 4)
        rng = np.random.default_rng(1234)
        self.base = (rng.standard_normal(1 << 24, dtype=np.float32) * np.float32(0.02))
        self.cache = {}

    def values(self, rows, k, salt):
        n = rows * k
        reps = (n + self.base.size - 1) // self.base.size
        off = (salt * 7919) % self.base.size
        src = np.roll(self.base, off)
        return (np.tile(src, reps)[:n] if reps > 1 else src[:n]).reshape(rows, k)

    def tensor(self, t, rows, k, salt):
        key = (t, rows, k, salt)
        if key in self.cache:
            return self.cache[key]
        x = self.values(rows, k, salt)
        if t == F32:
            out = x.copy()
        elif t == F16:
            out = x.astype(np.float16)
        else:
            rs = self.lib.ggml_row_size(t, k)
            out = np.empty((rows, rs), dtype=np.uint8)
            nchunk = min(rows, 4 * (os.cpu_count() or 4))
            bounds = [rows * i // nchunk for i in range(nchunk + 1)]

            def job(i):
                r0, r1 = bounds[i], bounds[i + 1]
                if r1 > r0:
                    xs = np.ascontiguousarray(x[r0:r1])
                    self.lib.ggml_quantize_chunk(t, xs.ctypes.data, out[r0:r1].ctypes.data, 0, r1 - r0, k, None)
            list(self.pool.map(job, range(nchunk)))
        self.cache[key] = out
        return out


def main():
    import gguf
    preset, ftype, path = sys.argv[1], sys.argv[2], sys.argv[3]
    n_embd, n_layer, n_head, n_kv, n_ff, n_vocab, n_expert, n_used, tied, n_ctx = PRESETS[preset]
    hd = n_embd // n_head
    is_70b = n_layer == 80
    q = Quantizer()
    w = gguf.GGUFWriter(path, "llama")
    w.add_name(preset + "-random-init")
    w.add_context_length(n_ctx)
    w.add_embedding_length(n_embd)
    w.add_block_count(n_layer)
    w.add_feed_forward_length(n_ff)
    w.add_head_count(n_head)
    w.add_head_count_kv(n_kv)
    w.add_rope_dimension_count(hd)
    w.add_rope_freq_base(500000.0 if "llama3" in preset else 1000000.0)
    w.add_layer_norm_rms_eps(1e-5)
    w.add_vocab_size(n_vocab)
    w.add_file_type(FTYPE_ID[ftype])
    if n_expert:
        w.add_expert_count(n_expert)
        w.add_expert_used_count(n_used)
    w.add_tokenizer_model("no_vocab")

    qt = {F32: gguf.GGMLQuantizationType.F32, F16: gguf.GGMLQuantizationType.F16, Q4_0: gguf.GGMLQuantizationType.Q4_0,
          Q8_0: gguf.GGMLQuantizationType.Q8_0, Q4_K: gguf.GGMLQuantizationType.Q4_K, Q5_K: gguf.GGMLQuantizationType.Q5_K,
          Q6_K: gguf.GGMLQuantizationType.Q6_K, Q4_1: gguf.GGMLQuantizationType.Q4_1, Q5_0: gguf.GGMLQuantizationType.Q5_0,
          Q5_1: gguf.GGMLQuantizationType.Q5_1, Q2_K: gguf.GGMLQuantizationType.Q2_K, Q3_K: gguf.GGMLQuantizationType.Q3_K,
          IQ4_NL: gguf.GGMLQuantizationType.IQ4_NL, IQ4_XS: gguf.GGMLQuantizationType.IQ4_XS}

    def add(full, short, il, rows, k, salt, experts=0):
        t = tensor_type(short, il, n_layer, ftype, n_expert, is_70b, tied)
        if short.endswith("_norm.weight"):
            w.add_tensor(full, np.ones(k, dtype=np.float32))
            return
        if experts:
            mats = [q.tensor(t, rows, k, salt + e) for e in range(experts)]
            data = np.stack(mats)
        else:
            data = q.tensor(t, rows, k, salt)
        if t in (F32, F16):
            w.add_tensor(full, data)
        else:
            w.add_tensor(full, data, raw_dtype=qt[t])

    add("token_embd.weight", "token_embd.weight", -1, n_vocab, n_embd, 1)
    for il in range(n_layer):
        p = "blk.%d." % il
        add(p + "attn_norm.weight", "attn_norm.weight", il, 1, n_embd, 0)
        add(p + "attn_q.weight", "attn_q.weight", il, n_embd, n_embd, 2)
        add(p + "attn_k.weight", "attn_k.weight", il, n_kv * hd, n_embd, 3)
        add(p + "attn_v.weight", "attn_v.weight", il, n_kv * hd, n_embd, 4)
        add(p + "attn_output.weight", "attn_output.weight", il, n_embd, n_embd, 5)
        add(p + "ffn_norm.weight", "ffn_norm.weight", il, 1, n_embd, 0)
        if n_expert:
            add(p + "ffn_gate_inp.weight", "ffn_gate_inp.weight", il, n_expert, n_embd, 6)
            add(p + "ffn_gate_exps.weight", "ffn_gate_exps.weight", il, n_ff, n_embd, 10, n_expert)
            add(p + "ffn_down_exps.weight", "ffn_down_exps.weight", il, n_embd, n_ff, 30, n_expert)
            add(p + "ffn_up_exps.weight", "ffn_up_exps.weight", il, n_ff, n_embd, 50, n_expert)
        else:
            add(p + "ffn_gate.weight", "ffn_gate.weight", il, n_ff, n_embd, 7)
            add(p + "ffn_down.weight", "ffn_down.weight", il, n_embd, n_ff, 8)
            add(p + "ffn_up.weight", "ffn_up.weight", il, n_ff, n_embd, 9)
    add("output_norm.weight", "output_norm.weight", -1, 1, n_embd, 0)
    if not tied:
        add("output.weight", "output.weight", -1, n_vocab, n_embd, 11)
    w.write_header_to_file()
    w.write_kv_data_to_file()
    w.write_tensors_to_file()
    w.close()
    print("wrote %s (%.2f GB)" % (path, os.path.getsize(path) / 1e9))


if __name__ == "__main__":
    main()
