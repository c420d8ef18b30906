"""CPU-only checks of the product's host side: the C-ABI library loads and exports every symbol
include/gpu_ec.h declares, the host algebra (coding / decode matrices, CRC combine, .meta
encoding, error text) matches the golden vectors and the oracle, and compute entry points fail
loudly — never fall back — when there is no GPU. No compute call needs a device here."""
import ctypes as C
import os
import re
import zlib

import numpy as np
import pytest

import gpu_ec
from gpu_ec import _lib, checksum, erasure
from gpu_ec.partition import count_for_rank, stripes_for_rank

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "gpu_ec.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(gpu_ec_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 30
    L = C.CDLL(_lib.SO_PATH)
    for name in sorted(declared):
        assert hasattr(L, name), f"{name} declared in gpu_ec.h but not exported"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)


def test_shard_len_reference_values():
    # erasure.rs:96-101 ; client/src/mod.rs:1553-1556
    assert erasure.shard_len(28, 4) == 7 and erasure.shard_len(10_000, 4) == 2500 and erasure.shard_len(1, 4) == 1
    assert erasure.shard_len(1000, 4) == 250 and erasure.shard_len(1001, 4) == 251 and erasure.shard_len(4, 4) == 1
    with pytest.raises(AssertionError):
        erasure.shard_len(10, 0)
    assert gpu_ec.lib().gpu_ec_shard_len(10, 0) == 0


def test_coding_matrix_matches_golden_and_oracle(golden, orc):
    for key, rows in golden["parity_rows"].items():
        k, m = map(int, key.split(","))
        M = erasure.coding_matrix(k, m)
        assert [bytes(r).hex() for r in M[k:]] == rows
        assert np.array_equal(M, orc.rs_matrix(k, m))
    for k, m in [(1, 1), (3, 2), (17, 3), (100, 28), (200, 56), (255, 1)]:
        assert np.array_equal(erasure.coding_matrix(k, m), orc.rs_matrix(k, m))


def test_decode_matrix_matches_golden_and_oracle(golden, orc):
    rows, miss, surv = erasure.decode_matrix(6, 3, [1, 0, 1, 1, 0, 1, 1, 0, 1])
    g = golden["decode_6_3_erase_1_4_7"]
    assert miss == [1, 4, 7] and surv == g["survivors"] and [bytes(r).hex() for r in rows] == g["rows"]
    rng = np.random.default_rng(3)
    for k, m in [(4, 2), (6, 3), (10, 4), (5, 5), (20, 6)]:
        for _ in range(10):
            present = np.ones(k + m, np.uint8)
            present[rng.choice(k + m, size=rng.integers(1, m + 1), replace=False)] = 0
            r1, m1, s1 = erasure.decode_matrix(k, m, present)
            r2, m2, s2 = orc.rs_decode_matrix(k, m, present)
            assert m1 == m2 and s1 == s2 and np.array_equal(r1, r2)
    with pytest.raises(gpu_ec.GpuEcError) as ei:
        erasure.decode_matrix(4, 2, [1, 0, 0, 0, 1, 1])
    assert ei.value.code == _lib.ERR_TOO_FEW_PRESENT and "TooFewShardsPresent" in str(ei.value)


def test_error_text_matches_reference():
    L = gpu_ec.lib()
    assert L.gpu_ec_strerror(_lib.ERR_ZERO_SHARDS) == b"data_shards and parity_shards must both be > 0"  # erasure.rs:9
    assert L.gpu_ec_strerror(_lib.ERR_EMPTY_DATA) == b"data must not be empty"  # erasure.rs:12
    assert L.gpu_ec_strerror(_lib.ERR_COUNT_MISMATCH) == b"Checksum count mismatch"  # chunkserver.rs:278
    assert L.gpu_ec_strerror(_lib.ERR_TOO_MANY_SHARDS).startswith(b"RS init error")
    assert L.gpu_ec_strerror(_lib.ERR_TOO_FEW_PRESENT).startswith(b"RS reconstruct error")


def test_argument_errors_precede_device_use():
    # these are decided before any CUDA call, so they hold with or without a GPU (erasure.rs:8-13)
    with pytest.raises(erasure.ErasureError, match="must both be > 0"):
        erasure.encode(b"abc", 0, 2)
    with pytest.raises(erasure.ErasureError, match="must both be > 0"):
        erasure.encode(b"abc", 4, 0)
    with pytest.raises(erasure.ErasureError, match="data must not be empty"):
        erasure.encode(b"", 4, 2)
    with pytest.raises(erasure.ErasureError, match="TooManyShards"):
        erasure.encode(b"abc", 200, 57)
    with pytest.raises(erasure.ErasureError, match="TooFewShardsPresent"):
        erasure.decode([np.zeros(4, np.uint8), None, None, None, None, np.zeros(4, np.uint8)], 4, 2, 16)
    with pytest.raises(erasure.ErasureError, match="IncorrectShardSize"):
        erasure.reconstruct([np.zeros(4, np.uint8), np.zeros(5, np.uint8), None, np.zeros(4, np.uint8)], 2, 2)
    # count mismatch is decided from the lengths alone (chunkserver.rs:277-279)
    if gpu_ec.lib().gpu_ec_device_count() > 0:
        assert checksum.verify_chunks(np.zeros(1025, np.uint8), b"\0" * 8) == "Checksum count mismatch"


def test_no_cpu_fallback_without_device():
    if gpu_ec.lib().gpu_ec_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(erasure.ErasureError, match="no usable CUDA device"):
        erasure.encode(b"Hello, Erasure Coding World!", 4, 2)
    with pytest.raises(gpu_ec.GpuEcError) as ei:
        checksum.crc32(b"123456789")
    assert ei.value.code == _lib.ERR_NO_DEVICE
    with pytest.raises(gpu_ec.GpuEcError):
        checksum.calculate_checksums(np.zeros(2048, np.uint8))


def test_meta_encode_and_combine(orc):
    crcs = np.array([0x3C7E92BB, 1, 0xFFFFFFFF], dtype=np.uint32)
    assert checksum.meta_bytes(crcs) == bytes.fromhex("3c7e92bb" "00000001" "ffffffff") == orc.meta_encode_be(crcs)
    rng = np.random.default_rng(5)
    d = rng.integers(0, 256, 70_000, dtype=np.uint8).tobytes()
    L = gpu_ec.lib()
    for cut in [0, 1, 511, 512, 65_536, 69_999, 70_000]:
        a, b = d[:cut], d[cut:]
        assert L.gpu_ec_crc32_combine(zlib.crc32(a), zlib.crc32(b), len(b), 0) == zlib.crc32(d)
        assert (L.gpu_ec_crc32_combine(orc.crc32(a, 1), orc.crc32(b, 1), len(b), 1) == orc.crc32(d, 1))


def test_round_robin_partition():
    for n, g in [(0, 1), (1, 8), (64, 1), (64, 8), (1000, 3), (4096, 8)]:
        seen = sorted(i for r in range(g) for i in stripes_for_rank(n, r, g))
        assert seen == list(range(n))
        assert [count_for_rank(n, r, g) for r in range(g)] == [len(stripes_for_rank(n, r, g)) for r in range(g)]
        assert max(count_for_rank(n, r, g) for r in range(g)) - min(count_for_rank(n, r, g) for r in range(g)) <= 1
