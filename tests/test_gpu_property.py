"""Property-based GPU parity (hypothesis): random (k, m, length, erasure pattern, chunk size) against
the CPU oracle and zlib, through the C ABI. Bit-exact: tolerance zero."""
import zlib

import numpy as np
import pytest
from hypothesis import HealthCheck, given, settings, strategies as st

pytestmark = pytest.mark.gpu

import gpu_ec  # noqa: E402
from gpu_ec import checksum, erasure  # noqa: E402

SET = settings(max_examples=60, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])


@st.composite
def code_and_data(draw):
    k = draw(st.integers(1, 20))
    m = draw(st.integers(1, 8))
    n = draw(st.one_of(st.integers(1, 4 * k), st.integers(1, 70_000), st.sampled_from([511, 512, 513, 8191, 8192, 8193, 65_536])))
    seed = draw(st.integers(0, 2**31 - 1))
    e = draw(st.integers(0, m))
    missing = draw(st.permutations(list(range(k + m))))[:e]
    return k, m, n, seed, sorted(missing)


@SET
@given(code_and_data())
def test_encode_decode_any_code(orc, case):
    k, m, n, seed, missing = case
    data = np.random.default_rng(seed).integers(0, 256, size=n, dtype=np.uint8)
    want = orc.encode(data, k, m)
    got, crcs, full, _ = erasure.encode_with_checksums(data, k, m)
    for i in range(k + m):
        assert np.array_equal(got[i], want[i]), (k, m, n, i)
        assert int(crcs[i]) == zlib.crc32(want[i].tobytes())
    assert full == zlib.crc32(data.tobytes())
    opt = [None if i in missing else want[i].copy() for i in range(k + m)]
    out = erasure.decode(opt, k, m, n)
    assert np.array_equal(out, data)
    for i in missing:
        assert np.array_equal(opt[i], want[i]), (k, m, missing, i)


@SET
@given(st.integers(0, 300_000), st.sampled_from([512, 128, 640, 1024, 2048, 4096, 65_536, 100, 7, 513]), st.integers(0, 2**31 - 1),
       st.sampled_from([0, 1]))
def test_crc_any_length_and_chunk(orc, n, chunk, seed, poly):
    d = np.random.default_rng(seed).integers(0, 256, size=n, dtype=np.uint8)
    assert checksum.crc32(d, poly) == orc.crc32(d, poly)
    got = checksum.calculate_checksums(d, chunk, poly)
    want = orc.crc32_chunks(d, chunk, poly) if n else np.zeros(0, np.uint32)
    assert np.array_equal(got, want)
    if n and poly == 0:
        meta = bytearray(orc.meta_encode_be(want))
        assert checksum.verify_chunks(d, bytes(meta), chunk) is None
        i = seed % len(want)
        meta[4 * i + (seed >> 8) % 4] ^= 1 << ((seed >> 12) % 8)
        assert checksum.verify_chunks(d, bytes(meta), chunk) == f"Checksum mismatch at chunk {i}"
        whole, chunks = checksum.block_checksums(d, chunk)
        assert whole == zlib.crc32(d.tobytes()) and np.array_equal(chunks, want)
