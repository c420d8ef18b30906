"""CPU checks of the sparse-fold CRC arithmetic (no GPU): the stride table compiled into the library satisfies
x^(128*Dv) == x^k0 + x^k1 + x^k2 + x^k3 (mod P_IEEE), and the word-for-word CPU model of the kernels + finisher
(tools/sfold_model.py) reproduces zlib.crc32 (the reference's crc32fast polynomial, chunkserver.rs:746-750)."""
import os
import re
import sys
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import sfold_model as sm  # noqa: E402


def _table():
    src = open(os.path.join(ROOT, "rust-hadoop-generated-by-llm_b200", "csrc", "sfold_kernels.cuh")).read()
    body = src[src.index("constexpr SfoldStride kSfoldIeee[]"):]
    body = body[:body.index("};")]
    return [(int(m.group(1)), [int(x) for x in m.group(2).split(",")]) for m in re.finditer(r"\{(\d+),\s*\{([\d,\s]+)\}\}", body)]


def test_stride_table_is_exact():
    tab = _table()
    assert len(tab) >= 6
    for dv, ks in tab:
        want = sm.xpow(128 * dv)
        got = 0
        for k in ks:
            got ^= sm.xpow(k)
        assert got == want, dv
        assert [k >> 5 for k in ks] == [0, 0, 1, 1] and all(k & 31 for k in ks)  # the shape sfold_mul is unrolled for


def test_model_matches_zlib_small_and_real_stride():
    rng = np.random.default_rng(7)
    for S, dv in [(4480, 40), (10_001, 33), (777, 70), (15, 5), (4097, 32)]:
        d = rng.integers(0, 256, S, dtype=np.uint8).tobytes()
        r = sm.xpow(128 * dv)
        ks = [31 - b for b in range(32) if (r >> b) & 1]   # any stride works with its dense polynomial
        assert sm.sfold_crc(d, dv, ks) == zlib.crc32(d)
    dv, ks = _table()[0]
    d = rng.integers(0, 256, 16 * dv * 3 - 5, dtype=np.uint8).tobytes()
    assert sm.sfold_crc(d, dv, ks) == zlib.crc32(d)
