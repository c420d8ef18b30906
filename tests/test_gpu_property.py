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


@st.composite
def fused_case(draw):
    k = draw(st.integers(1, 12))
    m = draw(st.integers(1, 4))
    # shards long enough for a sparse-fold stride (>= ~9 906 vectors) and enough stripes to fill the GPU
    S = draw(st.integers(160_000, 700_000))
    pad = draw(st.integers(0, k - 1))
    n_stripes = draw(st.integers(30, 48))
    seed = draw(st.integers(0, 2**31 - 1))
    return k, m, S * k - pad, n_stripes, seed


@settings(max_examples=10, deadline=None, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])
@given(fused_case())
def test_fused_encode_crc_any_code(orc, case):
    """gpu_ec_encode_crc_dev, single-pass path, random (k, m, length): parity + every shard CRC + CRC(original) of the
    first and last stripe against the oracle / zlib, and all stripes against the two-pass path."""
    import os
    import torch
    from gpu_ec import device
    k, m, stripe_len, n_stripes, seed = case
    S = erasure.shard_len(stripe_len, k)
    b = device.StripeBatch(n_stripes, k, m, S)
    b.fill_synthetic(seed, original_len=stripe_len)
    os.environ["GPU_EC_SFOLD"] = "0"
    crc_ref, full_ref = b.encode_crc(original_len=stripe_len)
    torch.cuda.synchronize()
    os.environ.pop("GPU_EC_SFOLD", None)
    crc, full = b.encode_crc(original_len=stripe_len)
    torch.cuda.synchronize()
    assert np.array_equal(device.u32(crc), device.u32(crc_ref)), case
    assert np.array_equal(device.u32(full), device.u32(full_ref)), case
    for i in (0, n_stripes - 1):
        data = np.concatenate([b.shard_host(i, j) for j in range(k)])[:stripe_len]
        want = orc.encode(data, k, m)
        assert all(np.array_equal(b.shard_host(i, s), want[s]) for s in range(k, k + m)), case
        assert device.u32(crc)[i].tolist() == [zlib.crc32(x.tobytes()) for x in want], case
        assert int(device.u32(full)[i]) == zlib.crc32(data.tobytes()), case
