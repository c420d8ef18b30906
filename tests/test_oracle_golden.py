"""The C oracle against the golden fixtures, the external fixed points and zlib (CPU only).

Mirrors the reference's own unit tests for the path: dfs/common/src/erasure.rs:61-109,
dfs/chunkserver/src/chunkserver.rs:1099-1140, dfs/client/src/mod.rs:1541-1557,1637-1664.
"""
import itertools
import zlib

import numpy as np
import pytest


def hx(s):
    return np.frombuffer(bytes.fromhex(s), dtype=np.uint8)


def test_field_spot_checks(orc, golden):
    assert orc.gf_mul(2, 128) == golden["field"]["2*128"] == 29
    assert orc.gf_mul(3, 7) == 9
    assert orc.gf_mul(0x53, 0xCA) == 143  # would be 1 under the AES polynomial 0x11B
    assert orc.gf_exp(0, 0) == 1 and orc.gf_exp(0, 3) == 0 and orc.gf_exp(2, 8) == 29


def test_parity_rows(orc, golden):
    for key, rows in golden["parity_rows"].items():
        k, m = map(int, key.split(","))
        M = orc.rs_matrix(k, m)
        assert np.array_equal(M[:k], np.eye(k, dtype=np.uint8))
        assert [bytes(r).hex() for r in M[k:]] == rows


def test_external_javareedsolomon_vector(orc, golden):
    g = golden["jrs_one_encode"]
    sh = orc.encode(hx(g["data"]), g["k"], g["m"])
    assert [bytes(s).hex() for s in sh[g["k"]:]] == g["parity"]


def test_hello_and_sevens(orc, golden):
    g = golden["hello_4_2"]
    sh = orc.encode(hx(g["data"]), 4, 2)
    assert [bytes(s).hex() for s in sh] == g["shards"]
    sh = orc.encode(np.full(4096, 7, dtype=np.uint8), 4, 2)
    assert (sh[4:] == 7).all()


def test_encode_cases(orc, golden):
    for c in golden["encode_cases"]:
        data = hx(c["data"])
        sh = orc.encode(data, c["k"], c["m"])
        assert [bytes(s).hex() for s in sh[c["k"]:]] == c["parity"]
        assert [orc.crc32(s) for s in sh] == c["shard_crc_ieee"]
        assert orc.crc32(data) == c["full_crc_ieee"]
        # systematic: first k shards concatenated + truncated == input (client/src/mod.rs:1637-1650)
        assert bytes(sh[:c["k"]].reshape(-1)[:data.size]) == bytes(data)


def test_shard_len(orc):
    # erasure.rs:96-101, client/src/mod.rs:1553-1556
    assert orc.shard_len(28, 4) == 7 and orc.shard_len(10_000, 4) == 2500 and orc.shard_len(1, 4) == 1
    assert orc.shard_len(1000, 4) == 250 and orc.shard_len(1001, 4) == 251 and orc.shard_len(4, 4) == 1
    assert orc.shard_len(64 << 20, 6) == 11_184_811 and orc.shard_len(1 << 20, 6) == 174_763


def test_encode_errors(orc):
    from oracle.oracle import OracleError
    with pytest.raises(OracleError):
        orc.encode(np.zeros(0, np.uint8), 4, 2)  # erasure.rs:103-108
    with pytest.raises(OracleError):
        orc.encode(np.zeros(8, np.uint8), 0, 2)
    with pytest.raises(OracleError):
        orc.encode(np.zeros(8, np.uint8), 4, 0)
    with pytest.raises(OracleError):
        orc.encode(np.zeros(8, np.uint8), 200, 57)


def test_decode_rows(orc, golden):
    for key, k, m, missing in [("decode_6_3_erase_1_4_7", 6, 3, [1, 4, 7]), ("decode_4_2_erase_1_4", 4, 2, [1, 4])]:
        present = [0 if i in missing else 1 for i in range(k + m)]
        rows, miss, surv = orc.rs_decode_matrix(k, m, present)
        assert miss == missing and surv == golden[key]["survivors"]
        assert [bytes(r).hex() for r in rows] == golden[key]["rows"]


@pytest.mark.parametrize("k,m", [(2, 2), (4, 2), (6, 3), (5, 5)])
def test_reconstruct_all_patterns(orc, k, m):
    rng = np.random.default_rng(k * 100 + m)
    data = rng.integers(0, 256, size=k * 37 - 3, dtype=np.uint8)
    sh = orc.encode(data, k, m)
    n = k + m
    for e in range(0, m + 1):
        for missing in itertools.combinations(range(n), e):
            present = np.ones(n, np.uint8)
            broken = sh.copy()
            for i in missing:
                present[i] = 0
                broken[i] = 0xAA
            fixed = orc.reconstruct(broken, present, k, m)
            assert np.array_equal(fixed, sh), (k, m, missing)
            out = orc.decode(broken, present, k, m, data.size)
            assert np.array_equal(out, data)
    from oracle.oracle import OracleError, ERR_TOO_FEW_PRESENT
    present = np.ones(n, np.uint8)
    present[: m + 1] = 0
    with pytest.raises(OracleError) as ei:
        orc.reconstruct(sh, present, k, m)
    assert ei.value.code == ERR_TOO_FEW_PRESENT


def test_reference_roundtrip_fixtures(orc):
    # erasure.rs:75-94 and client/src/mod.rs:1652-1664
    hello = np.frombuffer(b"Hello, Erasure Coding World!", dtype=np.uint8)
    sh = orc.encode(hello, 4, 2)
    present = np.array([1, 0, 1, 1, 0, 1], np.uint8)
    assert bytes(orc.decode(sh, present, 4, 2, hello.size)) == hello.tobytes()
    big = (np.arange(10_000) % 256).astype(np.uint8)
    sh = orc.encode(big, 4, 2)
    assert sh.shape == (6, 2500)
    assert np.array_equal(orc.decode(sh, np.ones(6, np.uint8), 4, 2, big.size), big)
    d = np.full(3333, 99, np.uint8)
    sh = orc.encode(d, 4, 2)
    assert np.array_equal(orc.decode(sh, np.array([1, 0, 1, 0, 1, 1], np.uint8), 4, 2, 3333), d)
    sh = orc.encode(np.zeros(1024, np.uint8), 4, 2)
    assert sh.shape == (6, 256)


def test_crc_check_values(orc, golden):
    chk = b"123456789"
    assert orc.crc32(chk) == golden["crc"]["external_check_ieee"] == zlib.crc32(chk)
    assert orc.crc32(chk, orc.POLY_CASTAGNOLI) == golden["crc"]["external_check_castagnoli"]
    blk = hx(golden["crc"]["block61"])
    assert orc.meta_encode_be(orc.crc32_chunks(blk)).hex() == golden["crc"]["block61_meta_be"]


def test_crc_cases(orc, golden):
    for c in golden["crc_cases"]:
        d = hx(c["data"])
        assert orc.crc32(d) == c["ieee"] == orc.crc32_bytewise(d)
        if c["castagnoli"] is not None:
            assert orc.crc32(d, orc.POLY_CASTAGNOLI) == c["castagnoli"]
        ch = orc.crc32_chunks(d)
        assert ch.tolist() == c["chunks512_ieee"]
        assert orc.meta_encode_be(ch).hex() == c["meta_be"]
        assert orc.verify_chunks(d, bytes.fromhex(c["meta_be"])) == (0, -1)


def test_crc_pclmul_path_vs_zlib(orc):
    rng = np.random.default_rng(7)
    for n in [64, 127, 128, 129, 143, 144, 191, 192, 255, 256, 1000, 4097, 1 << 16, (1 << 20) + 13]:
        d = rng.integers(0, 256, size=n, dtype=np.uint8)
        assert orc.crc32(d) == zlib.crc32(d.tobytes()), n
        assert orc.crc32(d, orc.POLY_CASTAGNOLI) == orc.crc32_bytewise(d, orc.POLY_CASTAGNOLI)


def test_crc_combine(orc):
    rng = np.random.default_rng(8)
    d = rng.integers(0, 256, size=5000, dtype=np.uint8)
    for cut in [0, 1, 512, 4999, 5000]:
        for poly in (orc.POLY_IEEE, orc.POLY_CASTAGNOLI):
            a, b = d[:cut], d[cut:]
            assert orc.crc32_combine(orc.crc32(a, poly), orc.crc32(b, poly), b.size, poly) == orc.crc32(d, poly)


def test_verify_semantics(orc):
    # chunkserver.rs:1117-1140: corrupt the first 9 bytes -> mismatch at chunk 0
    blk = np.frombuffer(b"Hello, world! This is a test block for checksum verification.", dtype=np.uint8).copy()
    meta = orc.meta_encode_be(orc.crc32_chunks(blk))
    assert orc.verify_chunks(blk, meta) == (0, -1)
    bad = blk.copy()
    bad[:9] = np.frombuffer(b"Corrupted", dtype=np.uint8)
    from oracle.oracle import ERR_COUNT_MISMATCH, ERR_CRC_MISMATCH
    assert orc.verify_chunks(bad, meta) == (ERR_CRC_MISMATCH, 0)
    big = (np.arange(65536) % 256).astype(np.uint8)
    meta = bytearray(orc.meta_encode_be(orc.crc32_chunks(big)))
    assert len(meta) == 4 * 128
    assert orc.verify_chunks(big, bytes(meta) + b"\x01\x02") == (0, -1)  # chunks_exact(4) drops the tail
    assert orc.verify_chunks(big[:-1], bytes(meta))[0] == ERR_CRC_MISMATCH
    assert orc.verify_chunks(big[:65536 - 512], bytes(meta))[0] == ERR_COUNT_MISMATCH
    meta[4 * 77 + 3] ^= 1
    assert orc.verify_chunks(big, bytes(meta)) == (ERR_CRC_MISMATCH, 77)
    assert orc.verify_chunks(np.zeros(0, np.uint8), b"") == (0, -1)


def test_synthetic_generator(orc, golden):
    g = golden["synthetic"]
    assert orc.fill_synthetic(64, g["seed"]).tobytes().hex() == g["first64"]
    assert orc.fill_synthetic(21, g["seed"], 1000003).tobytes().hex() == g["at_1000003_len_21"]


def test_oracle_against_reference_crates_when_pinned(orc):
    """tests/golden/ref_golden.json comes from the reference's OWN crates (tools/ref_harness, built with cargo).
    Neither this image nor the GPU boxes have a Rust toolchain (profiles/r02/r02_toolchain_probe.txt), so the file
    does not exist and parity stays 'unpinned' (oracle header); the day it exists this test pins it."""
    import json
    import os
    import zlib
    import numpy as np
    import pytest
    path = os.path.join(os.path.dirname(__file__), "golden", "ref_golden.json")
    if not os.path.exists(path):
        pytest.skip("no ref_golden.json: cargo is not available to build tools/ref_harness")
    ref = json.load(open(path))
    for c in ref["encode_cases"]:
        data = np.frombuffer(bytes.fromhex(c["data"]), dtype=np.uint8)
        got = orc.encode(data, c["k"], c["m"])
        assert [bytes(s).hex() for s in got] == c["shards"]
        assert [zlib.crc32(bytes(s)) for s in got] == c["shard_crc_ieee"]
