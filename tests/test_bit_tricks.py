"""CPU checks of the bit tricks the kernels rely on, with the constants read out of the CUDA sources (no GPU needed):
the PRMT table lookup of the IQ4 formats (csrc/common.cuh: iq4_lookup8), the multiply that spreads four qh bits over four bytes
(Q5_0 / Q5_1), and the "magic byte" bf16 construction of the MMQ dequantisers (csrc/mmq.cu)."""
import os
import re
import struct

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "llamacog_b200", "csrc")
IQ4_TABLE = [-127, -104, -83, -65, -49, -35, -22, -10, 1, 13, 25, 38, 53, 69, 89, 113]      # reference ggml-common.h:1077-1079


def _prmt(a, b, s):
    """byte_perm / PRMT: result byte i = byte (s >> 4i) & 7 of {a, b}; bit 3 of the selector nibble replicates the sign bit."""
    pool = [(a >> (8 * i)) & 0xFF for i in range(4)] + [(b >> (8 * i)) & 0xFF for i in range(4)]
    out = 0
    for i in range(4):
        sel = (s >> (4 * i)) & 0xF
        v = pool[sel & 7]
        if sel & 8:
            v = 0xFF if v & 0x80 else 0
        out |= v << (8 * i)
    return out


def test_iq4_table_words_in_the_source_hold_the_reference_table():
    src = open(os.path.join(CSRC, "common.cuh")).read()
    m = re.search(r"T0 = (0x[0-9a-fA-F]+)u, T1 = (0x[0-9a-fA-F]+)u, T2 = (0x[0-9a-fA-F]+)u, T3 = (0x[0-9a-fA-F]+)u", src)
    assert m, "iq4_lookup8 constants not found"
    words = [int(x, 16) for x in m.groups()]
    got = list(struct.unpack("<16b", struct.pack("<4I", *words)))
    assert got == IQ4_TABLE
    dq = open(os.path.join(CSRC, "dequant.cuh")).read()
    m = re.search(r"iq4_values\[16\] = \{([^}]*)\}", dq)
    assert [int(x) for x in m.group(1).split(",")] == IQ4_TABLE


def test_iq4_lookup_by_three_picks():
    src = open(os.path.join(CSRC, "common.cuh")).read()
    T = [int(x, 16) for x in re.search(r"T0 = (0x\w+)u, T1 = (0x\w+)u, T2 = (0x\w+)u, T3 = (0x\w+)u", src).groups()]
    rng = np.random.default_rng(0)
    for q in [int(x) for x in rng.integers(0, 1 << 32, 4000, dtype=np.uint64)] + [0, 0xFFFFFFFF, 0x0F0F0F0F, 0xF0F0F0F0, 0x76543210, 0xFEDCBA98]:
        r = []
        for i in range(2):
            idx = (q >> (16 * i)) & 0xFFFF
            a = _prmt(T[0], T[1], idx & 0x7777)
            b = _prmt(T[2], T[3], idx & 0x7777)
            r.append(_prmt(a, b, 0x3210 | ((idx & 0x8888) >> 1)))
        lo, hi = _prmt(r[0], r[1], 0x6420), _prmt(r[0], r[1], 0x7531)
        for k in range(4):
            byte = (q >> (8 * k)) & 0xFF
            assert struct.unpack("<b", bytes([(lo >> (8 * k)) & 0xFF]))[0] == IQ4_TABLE[byte & 0xF]
            assert struct.unpack("<b", bytes([(hi >> (8 * k)) & 0xFF]))[0] == IQ4_TABLE[byte >> 4]


def test_spread_four_bits_over_four_bytes():
    """x * 0x00204081 puts bit k of a 4-bit x at bit 8k (the four shifted copies do not overlap): the fifth bits of Q5_0 / Q5_1."""
    for x in range(16):
        v = ((x * 0x00204081) & 0x01010101)
        assert [(v >> (8 * k)) & 0xFF for k in range(4)] == [(x >> k) & 1 for k in range(4)]
    for f in ("mmvq_units.cuh", "mmvq_r.cuh", "mmq.cu"):
        assert "0x00204081u" in open(os.path.join(CSRC, f)).read(), f


def _bf16_to_float(h):
    return struct.unpack("<f", struct.pack("<I", h << 16))[0]


def test_magic_byte_is_one_plus_code_over_two_to_the_n():
    """A byte 0x80 | (q << s) under a 0x3F high byte is the bf16 value 1 + q / 2^(7-s); subtracting 1.5 centres the code exactly."""
    for s, bits in ((3, 4), (2, 5), (1, 6), (5, 2), (4, 3)):           # Q4_*, Q5_*, Q6_K, Q2_K, Q3_K code positions
        n = 7 - s
        for q in range(1 << bits):
            h = 0x3F00 | 0x80 | (q << s)
            assert h <= 0x3FFF
            v = _bf16_to_float(h)
            assert v == 1.0 + q / 2.0 ** n
            centred = np.float32(v) - np.float32(1.5)
            assert float(centred) == (q - 2 ** (n - 1)) / 2.0 ** n          # exact in bf16: at most 7 significant bits
    assert _bf16_to_float(0xBFC0) == -1.5                                # the M15 constant of expand4 / expand4_sym


def test_f16_bias_trick_for_int8_codes():
    """1024 + (v ^ 0x80) is exact in f16 (0x6400 | byte), minus 1152 gives v: Q8_0 and the IQ4 table values in the MMQ."""
    for v in range(-128, 128):
        byte = (v & 0xFF) ^ 0x80
        h = np.array([0x6400 | byte], dtype=np.uint16).view(np.float16)[0]
        assert float(h) == 1024 + byte and float(h) - 1152.0 == v
