"""GPU parity: the CUDA path, called through the C ABI of libgpu_ec.so, against the CPU oracle,
the committed golden fixtures and zlib — bit-exact (byte / integer work: tolerance is zero).

Cases mirror the reference's own tests (dfs/common/src/erasure.rs:61-109,
dfs/chunkserver/src/chunkserver.rs:1099-1198, dfs/client/src/mod.rs:1541-1664,
test_scripts/erasure_coding_test.sh:66-153) plus the edge cases of SURVEY.md §7 H3/H4.
"""
import ctypes as C
import itertools
import threading
import zlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

import gpu_ec  # noqa: E402
from gpu_ec import _lib, checksum, erasure  # noqa: E402


def hx(s):
    return np.frombuffer(bytes.fromhex(s), dtype=np.uint8)


@pytest.fixture(scope="module")
def L():
    lib = gpu_ec.lib()
    assert lib.gpu_ec_device_count() > 0, "GPU tests need a CUDA device"
    gpu_ec._lib.check(lib.gpu_ec_init())
    return lib


def rnd(n, seed):
    return np.random.default_rng(seed).integers(0, 256, size=n, dtype=np.uint8)


# ------------------------------------------------------------------ encode
def test_encode_golden_cases(L, golden, orc):
    before = L.gpu_ec_kernel_launch_count()
    for c in golden["encode_cases"]:
        data = hx(c["data"])
        k, m = c["k"], c["m"]
        shards, crcs, full, metas = erasure.encode_with_checksums(data, k, m, with_meta=True)
        assert [bytes(s).hex() for s in shards[k:]] == c["parity"]
        assert crcs.tolist() == c["shard_crc_ieee"]
        assert full == c["full_crc_ieee"]
        assert bytes(np.concatenate(shards[:k])[:data.size]) == bytes(data)
        for s, mt in zip(shards, metas):
            assert mt.tobytes() == orc.meta_encode_be(orc.crc32_chunks(s))
    assert L.gpu_ec_kernel_launch_count() > before  # the CUDA kernels really ran


def test_encode_external_and_reference_fixtures(L, golden):
    g = golden["jrs_one_encode"]
    sh = erasure.encode(hx(g["data"]), g["k"], g["m"])
    assert [bytes(s).hex() for s in sh[g["k"]:]] == g["parity"]
    g = golden["hello_4_2"]
    assert [bytes(s).hex() for s in erasure.encode(hx(g["data"]), 4, 2)] == g["shards"]
    sh = erasure.encode(np.full(4096, 7, np.uint8), 4, 2)  # chunkserver.rs:1102-1103
    assert all((s == 7).all() for s in sh)
    sh = erasure.encode(np.zeros(1024, np.uint8), 4, 2)  # client/src/mod.rs:1541-1551
    assert len(sh) == 6 and all(s.size == 256 for s in sh)


@pytest.mark.parametrize("k,m", [(2, 2), (4, 2), (5, 5), (6, 3), (10, 4), (1, 1), (7, 4), (3, 1)])
def test_encode_sizes_vs_oracle(L, orc, k, m):
    for n in [1, 2, 15, 28, 511, 512, 513, 4096, 10_000, 65_536, (1 << 20) + 7]:
        data = rnd(n, n * 31 + k)
        want = orc.encode(data, k, m)
        got = erasure.encode(data, k, m)
        assert len(got) == k + m
        for i in range(k + m):
            assert np.array_equal(got[i], want[i]), (k, m, n, i)


@pytest.mark.parametrize("k,m", [(12, 4), (8, 8), (9, 3), (11, 2), (13, 3), (14, 2), (20, 6), (56, 4), (60, 10), (100, 28), (250, 6)])
def test_encode_wide_codes_vs_oracle(L, orc, k, m):
    # 8 <= k <= 12: 2-CTA cluster kernel (odd k: uneven split); 13, 14: two byte-LUT passes;
    # beyond: split-nibble kernel; m > 4: several output passes; k > 56: accumulate passes
    for n in [k * 3 + 1, 50_000, 3_000_017]:
        data = rnd(n, k * 1000 + m)
        want = orc.encode(data, k, m)
        got = erasure.encode(data, k, m)
        for i in range(k + m):
            assert np.array_equal(got[i], want[i]), (k, m, n, i)


def test_encode_adversarial_fills(L, orc):
    n = 200_003
    for fill in [np.zeros(n, np.uint8), np.full(n, 0xFF, np.uint8), (np.arange(n) % 256).astype(np.uint8)]:
        for k, m in [(6, 3), (10, 4)]:
            want = orc.encode(fill, k, m)
            got = erasure.encode(fill, k, m)
            assert all(np.array_equal(a, b) for a, b in zip(got, want))


def test_encode_config1_64mib_block(L, orc):
    # BASELINE config 1: RS(6,3) over a 64 MiB block, S = 11 184 811 (odd)
    n = 64 << 20
    data = orc.fill_synthetic(n, 0xEC000001)
    want = orc.encode(data, 6, 3)
    shards, crcs, full, _ = erasure.encode_with_checksums(data, 6, 3)
    assert shards[0].size == 11_184_811
    for i in range(9):
        assert np.array_equal(shards[i], want[i]), i
        assert int(crcs[i]) == zlib.crc32(want[i].tobytes())
    assert full == zlib.crc32(data.tobytes())


def test_encode_errors(L):
    with pytest.raises(erasure.ErasureError, match="data must not be empty"):
        erasure.encode(b"", 4, 2)  # erasure.rs:103-108
    with pytest.raises(erasure.ErasureError, match="must both be > 0"):
        erasure.encode(b"x", 0, 1)
    with pytest.raises(erasure.ErasureError, match="TooManyShards"):
        erasure.encode(b"x" * 1000, 255, 2)


# ------------------------------------------------------------------ reconstruct / decode
@pytest.mark.parametrize("k,m", [(2, 2), (4, 2), (6, 3)])
def test_reconstruct_all_patterns(L, orc, k, m):
    data = rnd(k * 1237 - 5, k + 17 * m)
    sh = orc.encode(data, k, m)
    n = k + m
    for e in range(0, m + 1):
        for missing in itertools.combinations(range(n), e):
            opt = [None if i in missing else sh[i].copy() for i in range(n)]
            erasure.reconstruct(opt, k, m)
            for i in range(n):
                assert np.array_equal(opt[i], sh[i]), (k, m, missing, i)
    opt = [None] * (m + 1) + [sh[i].copy() for i in range(m + 1, n)]
    with pytest.raises(erasure.ErasureError, match="TooFewShardsPresent"):
        erasure.reconstruct(opt, k, m)


@pytest.mark.parametrize("k,m", [(5, 5), (10, 4), (20, 6), (60, 10)])
def test_reconstruct_random_patterns_wide(L, orc, k, m):
    rng = np.random.default_rng(k * m)
    data = rnd(k * 301 + 3, k)
    sh = orc.encode(data, k, m)
    n = k + m
    for _ in range(6):
        missing = set(rng.choice(n, size=m, replace=False).tolist())
        opt = [None if i in missing else sh[i].copy() for i in range(n)]
        out = erasure.decode(opt, k, m, data.size)
        assert np.array_equal(out, data)
        for i in range(n):
            assert np.array_equal(opt[i], sh[i])


def test_reference_roundtrip_fixtures(L):
    hello = np.frombuffer(b"Hello, Erasure Coding World!", dtype=np.uint8)
    sh = erasure.encode(hello, 4, 2)  # erasure.rs:66-84
    opt = [s.copy() for s in sh]
    assert bytes(erasure.decode(opt, 4, 2, hello.size)) == hello.tobytes()
    opt = [s.copy() for s in sh]
    opt[1] = None
    opt[4] = None
    assert bytes(erasure.decode(opt, 4, 2, hello.size)) == hello.tobytes()
    assert np.array_equal(opt[1], sh[1]) and np.array_equal(opt[4], sh[4])
    big = (np.arange(10_000) % 256).astype(np.uint8)  # erasure.rs:86-94
    sh = erasure.encode(big, 4, 2)
    assert np.array_equal(erasure.decode([s.copy() for s in sh], 4, 2, big.size), big)
    d = np.full(4096, 7, np.uint8)  # chunkserver.rs:1099-1115
    sh = erasure.encode(d, 4, 2)
    opt = [s.copy() for s in sh]
    opt[4] = None
    erasure.reconstruct(opt, 4, 2)
    assert np.array_equal(opt[4], sh[4])
    d = np.full(5000, 42, np.uint8)  # client/src/mod.rs:1637-1650 fast path
    sh = erasure.encode(d, 4, 2)
    assert np.array_equal(np.concatenate(sh[:4])[:5000], d)
    d = np.full(3333, 99, np.uint8)  # client/src/mod.rs:1652-1664
    sh = erasure.encode(d, 4, 2)
    opt = [s.copy() for s in sh]
    opt[1] = None
    opt[3] = None
    assert np.array_equal(erasure.decode(opt, 4, 2, 3333), d)
    blk = np.tile(np.arange(256, dtype=np.uint8), 256)  # erasure_coding_test.sh:66-153, RS(2,2) 64 KiB
    sh = erasure.encode(blk, 2, 2)
    opt = [None, sh[1].copy(), sh[2].copy(), None]
    assert np.array_equal(erasure.decode(opt, 2, 2, blk.size), blk)


def test_reconstruct_large_single_call_chunked(L, orc):
    """A >= 64 MiB stripe through the host-pointer reconstruct: the call walks it in column chunks with the rebuilt
    shards copied back on a second stream (chunkserver.rs:620-623 on a large block); bytes must not change."""
    k, m = 6, 3
    data = orc.fill_synthetic((66 << 20) + 13, 0xEC0000E1)
    sh = orc.encode(data, k, m)
    for missing in ([1, 4, 7], [0], [6, 8]):
        opt = [None if i in missing else sh[i].copy() for i in range(k + m)]
        erasure.reconstruct(opt, k, m)
        for i in missing:
            assert np.array_equal(opt[i], sh[i]), (missing, i)


def test_reconstruct_config3_worst_case_stripe(L, orc):
    # one 1 MiB stripe at RS(6,3), 2 data + 1 parity lost: S = 174 763
    data = orc.fill_synthetic(1 << 20, 0xEC000003)
    sh = orc.encode(data, 6, 3)
    assert sh.shape[1] == 174_763
    opt = [None if i in (1, 4, 7) else sh[i].copy() for i in range(9)]
    erasure.reconstruct(opt, 6, 3)
    for i in range(9):
        assert np.array_equal(opt[i], sh[i])


# ------------------------------------------------------------------ CRC
def test_crc_golden_and_check_values(L, golden):
    assert checksum.crc32(b"123456789") == 0xCBF43926
    assert checksum.crc32(b"123456789", gpu_ec.POLY_CASTAGNOLI) == 0xE3069283
    assert checksum.crc32(b"") == 0
    blk = hx(golden["crc"]["block61"])
    assert checksum.meta_bytes(checksum.calculate_checksums(blk)).hex() == golden["crc"]["block61_meta_be"]
    for c in golden["crc_cases"]:
        d = hx(c["data"])
        assert checksum.crc32(d) == c["ieee"]
        if c["castagnoli"] is not None:
            assert checksum.crc32(d, gpu_ec.POLY_CASTAGNOLI) == c["castagnoli"]
        ch = checksum.calculate_checksums(d)
        assert ch.tolist() == c["chunks512_ieee"]
        assert checksum.meta_bytes(ch).hex() == c["meta_be"]
        assert checksum.verify_chunks(d, bytes.fromhex(c["meta_be"])) is None


@pytest.mark.parametrize("n", [1, 3, 16, 127, 128, 129, 4095, 4096, 4097, 16_383, 16_384, 16_385, 174_763,
                               (1 << 20) + 511, 11_184_811, (32 << 20) + 1])
def test_crc_whole_vs_zlib(L, orc, n):
    d = rnd(n, n)
    assert checksum.crc32(d) == zlib.crc32(d.tobytes())
    if n <= 1 << 21:
        assert checksum.crc32(d, gpu_ec.POLY_CASTAGNOLI) == orc.crc32(d, orc.POLY_CASTAGNOLI)


@pytest.mark.parametrize("chunk", [512, 65_536, 128, 640, 1024, 4096, 8192, 1000, 7, 1 << 20])
def test_crc_chunks_vs_oracle(L, orc, chunk):
    for n in [chunk - 1, chunk, chunk + 1, 5 * chunk + 171, 300_000, (8 << 20) + 171]:
        if n <= 0 or (chunk < 128 and n > 100_000):
            continue
        d = rnd(n, n + chunk)
        got = checksum.calculate_checksums(d, chunk)
        want = orc.crc32_chunks(d, chunk)
        assert np.array_equal(got, want), (chunk, n, np.nonzero(got != want)[0][:4])
    d = rnd(200_000, chunk)
    assert np.array_equal(checksum.calculate_checksums(d, chunk, gpu_ec.POLY_CASTAGNOLI),
                          orc.crc32_chunks(d, chunk, orc.POLY_CASTAGNOLI))


def test_crc_whole_multi_gib_three_level_fold(L):
    """> 2^20 segments in one buffer: the fold runs three levels (1024^2 < segments)."""
    import torch
    from gpu_ec import device
    n = (1 << 29) + 512 * 1025 + 77  # > 2^20 512-byte segments, ragged tail
    t = torch.empty(n, dtype=torch.uint8, device="cuda")
    device.fill_synthetic(t, 0xEC0000C3)
    # force 512-byte segments through the chunk path: 2^20+ chunk CRCs, then compare their fold with the whole CRC
    chunks = device.crc32_chunks(t, 512)
    whole = int(device.u32(device.crc32(t))[0])
    host = t.cpu().numpy()
    assert whole == zlib.crc32(host.tobytes())
    ch = device.u32(chunks)
    for i in (0, 1, 1 << 19, ch.size - 2, ch.size - 1):
        assert int(ch[i]) == zlib.crc32(host[i * 512:(i + 1) * 512].tobytes())
    w2, c2 = checksum.block_checksums(host)  # host path: whole CRC folded from 2^20+ chunk CRCs (3 levels)
    assert w2 == whole and np.array_equal(c2, ch)


def test_block_checksums_one_pass(L, orc):
    for n in [61, 512, 513, 65_536, 11_184_811]:
        d = rnd(n, 99 + n)
        whole, chunks = checksum.block_checksums(d)
        assert whole == zlib.crc32(d.tobytes())
        assert np.array_equal(chunks, orc.crc32_chunks(d))
    whole, chunks = checksum.block_checksums(rnd(300_000, 5), chunk=65_536)
    assert whole == zlib.crc32(rnd(300_000, 5).tobytes()) and chunks.size == 5


def test_block_checksums_batch_one_launch(L, orc):
    """Many WriteBlocks at once (chunkserver.rs:746-769 + :195 each): same CRCs as block by block, as the oracle,
    as zlib; ragged last chunks, empty blocks, and exactly one CRC launch + one fold launch for the whole batch."""
    sizes = [0, 1, 61, 511, 512, 513, 4096, 65_536, 65_537, 1 << 20, 1_000_003, 0, 700, 131_072 + 17]
    blocks = [rnd(n, 900 + i) for i, n in enumerate(sizes)]
    before = L.gpu_ec_kernel_launch_count()
    whole, chunks = checksum.block_checksums_batch(blocks)
    assert L.gpu_ec_kernel_launch_count() - before == 2
    for b, w, c in zip(blocks, whole, chunks):
        assert int(w) == zlib.crc32(b.tobytes()), b.size
        assert np.array_equal(c, orc.crc32_chunks(b)), b.size
        w1, c1 = checksum.block_checksums(b)
        assert w1 == int(w) and np.array_equal(c1, c)
    # Castagnoli and another chunk size; a chunk size the one-launch layout does not serve falls back block by block
    whole, chunks = checksum.block_checksums_batch(blocks, chunk=1024, poly=gpu_ec.POLY_CASTAGNOLI)
    for b, w, c in zip(blocks, whole, chunks):
        assert int(w) == orc.crc32(b, orc.POLY_CASTAGNOLI) and np.array_equal(c, orc.crc32_chunks(b, 1024, orc.POLY_CASTAGNOLI))
    whole, chunks = checksum.block_checksums_batch(blocks[:6], chunk=4096)
    for b, w, c in zip(blocks[:6], whole, chunks):
        assert int(w) == zlib.crc32(b.tobytes()) and np.array_equal(c, orc.crc32_chunks(b, 4096))
    assert checksum.block_checksums_batch([])[1] == []
    big = [rnd((1 << 20) + 300 * i, 70 + i) for i in range(70)]  # > 64 MiB in total: copied block by block, not packed
    whole, chunks = checksum.block_checksums_batch(big)
    for b, w, c in zip(big[::23], whole[::23], chunks[::23]):
        assert int(w) == zlib.crc32(b.tobytes()) and np.array_equal(c, orc.crc32_chunks(b))


def test_verify_semantics(L, orc):
    blk = np.frombuffer(b"Hello, world! This is a test block for checksum verification.", dtype=np.uint8).copy()
    meta = checksum.meta_bytes(checksum.calculate_checksums(blk))
    assert checksum.verify_chunks(blk, meta) is None
    bad = blk.copy()
    bad[:9] = np.frombuffer(b"Corrupted", dtype=np.uint8)  # chunkserver.rs:1128-1139
    assert checksum.verify_chunks(bad, meta) == "Checksum mismatch at chunk 0"
    big = (np.arange(65_536) % 256).astype(np.uint8)
    meta = bytearray(checksum.meta_bytes(checksum.calculate_checksums(big)))
    assert len(meta) == 512
    assert checksum.verify_chunks(big, bytes(meta) + b"\x01\x02") is None  # chunks_exact(4) ignores the tail
    assert checksum.verify_chunks(big[:-1], bytes(meta)) == "Checksum mismatch at chunk 127"
    assert checksum.verify_chunks(big[:65_536 - 512], bytes(meta)) == "Checksum count mismatch"
    meta[4 * 77 + 3] ^= 1
    meta[4 * 100] ^= 0x80
    assert checksum.verify_chunks(big, bytes(meta)) == "Checksum mismatch at chunk 77"  # FIRST bad chunk
    assert checksum.verify_chunks(np.zeros(0, np.uint8), b"") is None
    assert checksum.check_expected(blk, 0) is None  # 0 = skip (chunkserver.rs:746)
    assert checksum.check_expected(blk, zlib.crc32(blk.tobytes())) is None
    msg = checksum.check_expected(blk, 12345)
    assert msg == f"Checksum mismatch: expected 12345, actual {zlib.crc32(blk.tobytes())}"
    # a large block with one flipped bit deep inside
    d = rnd((16 << 20) + 100, 1)
    meta = checksum.meta_bytes(checksum.calculate_checksums(d))
    d[12_345_678] ^= 4
    assert checksum.verify_chunks(d, meta) == f"Checksum mismatch at chunk {12_345_678 // 512}"


def test_block_store_mirrors_chunkserver_tests(L, tmp_path):
    store = checksum.BlockStore(str(tmp_path))
    data = b"Hello, world! This is a test block for checksum verification."
    store.write_block("test_block", data)  # chunkserver.rs:1117-1140
    store.verify_block("test_block", data)
    assert (tmp_path / "test_block.meta").read_bytes().hex() == "3c7e92bb"
    corrupted = b"Corrupted" + data[9:]
    with pytest.raises(ValueError, match="Checksum mismatch at chunk 0"):
        store.verify_block("test_block", corrupted)
    with pytest.raises(ValueError, match="Checksum file missing"):
        store.verify_block("nope", data)
    blk = (np.arange(64 * 1024) % 256).astype(np.uint8)  # chunkserver.rs:1146-1198
    store.write_block("b64k", blk)
    assert (tmp_path / "b64k.meta").stat().st_size == 4 * 128
    store.verify_partial_read("b64k", 32_768, 4096)
    store.verify_partial_read("b64k", 0, 1)
    store.verify_partial_read("b64k", 65_535, 1)
    raw = bytearray((tmp_path / "b64k").read_bytes())
    raw[33_000] ^= 1
    (tmp_path / "b64k").write_bytes(bytes(raw))
    with pytest.raises(ValueError, match=f"Checksum mismatch at chunk {33_000 // 512}"):
        store.verify_partial_read("b64k", 32_768, 4096)
    store.verify_partial_read("b64k", 0, 32_768)  # untouched range still verifies
    store.write_block("empty", b"")
    assert (tmp_path / "empty.meta").stat().st_size == 0
    store.verify_block("empty", b"")


# ------------------------------------------------------------------ device-resident entry points
def test_synthetic_generator_matches_oracle(L, orc, golden):
    import torch
    from gpu_ec import device
    t = torch.empty(100_003, dtype=torch.uint8, device="cuda")
    device.fill_synthetic(t, golden["synthetic"]["seed"], 0)
    host = t.cpu().numpy()
    assert host[:64].tobytes().hex() == golden["synthetic"]["first64"]
    assert np.array_equal(host, orc.fill_synthetic(100_003, golden["synthetic"]["seed"]))
    device.fill_synthetic(t[:4096], 77, 8 * 12_345)
    assert np.array_equal(t[:4096].cpu().numpy(), orc.fill_synthetic(4096, 77, 8 * 12_345))


@pytest.mark.parametrize("k,m,stripe_len,n_stripes", [(6, 3, 1 << 20, 24), (10, 4, (1 << 20) + 3, 9), (4, 2, 4099, 300),
                                                       (6, 3, 5, 7)])
def test_dev_encode_reconstruct_crc(L, orc, k, m, stripe_len, n_stripes):
    import torch
    from gpu_ec import device
    S = erasure.shard_len(stripe_len, k)
    b = device.StripeBatch(n_stripes, k, m, S)
    b.fill_synthetic(0xEC0000AA, original_len=stripe_len)
    torch.cuda.synchronize()
    data0 = [b.shard_host(0, j) for j in range(k)]
    assert np.array_equal(data0[1], orc.fill_synthetic(S, 0xEC0000AA, b.stream_offset(0, 1)))
    shard_crc, full = b.encode_crc(original_len=stripe_len)
    torch.cuda.synchronize()
    shard_crc, full = device.u32(shard_crc), device.u32(full)
    M = orc.rs_matrix(k, m)
    for i in sorted({0, 1, n_stripes // 2, n_stripes - 1}):
        data = np.concatenate([b.shard_host(i, j) for j in range(k)])
        want = orc.encode(data, k, m)
        for s in range(k + m):
            got = b.shard_host(i, s)
            assert np.array_equal(got, want[s]), (i, s)
            assert int(shard_crc[i, s]) == zlib.crc32(want[s].tobytes()), (i, s)
        assert int(full[i]) == zlib.crc32(data[:stripe_len].tobytes())
    assert M.shape == (k + m, k)
    assert np.array_equal(device.u32(b.shard_crcs()), shard_crc)
    # erase m shards (worst case mixes data and parity) in every stripe, rebuild in one pass
    golden_buf = b.buf.clone()
    missing = [1, k - 1, k + m - 1][:m] if m >= 3 else ([1, k + m - 1][:m])
    present = np.ones(k + m, np.uint8)
    for s in missing:
        present[s] = 0
        b.buf[:, s, :] = 0x5A
    b.reconstruct(present)
    torch.cuda.synchronize()
    assert torch.equal(b.buf[:, :, :S], golden_buf[:, :, :S])


def test_dev_crc_chunks_and_verify(L, orc):
    import torch
    from gpu_ec import device
    n = (48 << 20) + 171
    t = torch.empty(n, dtype=torch.uint8, device="cuda")
    device.fill_synthetic(t, 0xEC000002)
    host = t.cpu().numpy()
    for chunk in (512, 65_536):
        got = device.u32(device.crc32_chunks(t, chunk))
        assert np.array_equal(got, orc.crc32_chunks(host, chunk)), chunk
    assert int(device.u32(device.crc32(t))[0]) == zlib.crc32(host.tobytes())
    assert int(device.u32(device.crc32(t, gpu_ec.POLY_CASTAGNOLI))[0]) == orc.crc32(host, orc.POLY_CASTAGNOLI)
    meta = device.crc32_chunks(t, 512, big_endian=True)
    assert device.u32(meta).tobytes() == orc.meta_encode_be(orc.crc32_chunks(host, 512))
    meta_u8 = meta.view(torch.uint8)
    assert int(device.verify_chunks(t, 512, meta_u8).item()) == -1
    t[40_000_000] ^= 1
    t[10_000_000] ^= 1
    assert int(device.verify_chunks(t, 512, meta_u8).item()) == 10_000_000 // 512


def test_dev_crc_chunks_many_buffers_with_fold(L, orc):
    """Several pitched buffers, chunk sizes that need the fold (chunk > segment), ragged tails, both byte orders."""
    import torch
    n_buf, pitch = 5, 300_032
    t = torch.empty(n_buf * pitch, dtype=torch.uint8, device="cuda")
    from gpu_ec import device
    device.fill_synthetic(t, 0xEC0000B7)
    host = t.cpu().numpy().reshape(n_buf, pitch)
    st = torch.cuda.current_stream().cuda_stream
    for length, chunk in [(300_000, 4096), (299_999, 65_536), (8191, 2048), (262_144, 131_072), (300_000, 1536)]:
        nch = (length + chunk - 1) // chunk
        for be in (0, 1):
            out = torch.zeros((n_buf, nch + 3), dtype=torch.int32, device="cuda")
            _lib.check(L.gpu_ec_crc32_chunks_dev(t.data_ptr(), n_buf, pitch, length, chunk, 0, out.data_ptr(), nch + 3, be, st))
            got = device.u32(out)
            for b in range(n_buf):
                want = orc.crc32_chunks(host[b, :length], chunk)
                if be:
                    want = want.byteswap()
                assert np.array_equal(got[b, :nch], want), (length, chunk, be, b)
                assert not got[b, nch:].any()  # nothing written past a buffer's own CRCs


def test_dev_crc_unaligned_views(L, orc):
    """Device buffers that are not 16-byte aligned take the generic (still CUDA) kernel."""
    import torch
    from gpu_ec import device
    t = torch.empty(300_000, dtype=torch.uint8, device="cuda")
    device.fill_synthetic(t, 0xEC0000A1)
    host = t.cpu().numpy()
    for off in (1, 3, 8, 15):
        v = t[off:off + 200_001]
        assert int(device.u32(device.crc32(v))[0]) == zlib.crc32(host[off:off + 200_001].tobytes())
        assert np.array_equal(device.u32(device.crc32_chunks(v, 512)), orc.crc32_chunks(host[off:off + 200_001], 512))


def test_linearity_and_checksum_of_checksums_full_size(L):
    """Size-independent properties at BASELINE sizes (no oracle pass over the whole input)."""
    import torch
    from gpu_ec import device
    k, m, S = 6, 3, 11_184_811
    a = device.StripeBatch(2, k, m, S)
    a.fill_synthetic(1)
    a.encode()
    x = a.buf[0, :k] ^ a.buf[1, :k]
    c = device.StripeBatch(1, k, m, S)
    c.buf[0, :k] = x
    c.encode()
    assert torch.equal(c.buf[0, k:, :S], (a.buf[0, k:] ^ a.buf[1, k:])[:, :S])  # encode(a^b) == encode(a)^encode(b)
    # erase -> reconstruct round trip at full size
    ref = a.buf.clone()
    a.buf[:, 0] = 0
    a.buf[:, 3] = 0
    a.buf[:, 8] = 0
    a.reconstruct([0, 1, 1, 0, 1, 1, 1, 1, 0])
    assert torch.equal(a.buf[:, :, :S], ref[:, :, :S])
    # whole-buffer CRC == fold of the 512-byte chunk CRCs (checksum of checksums)
    t = a.buf[0, 0, :S].contiguous()
    chunks = device.u32(device.crc32_chunks(t, 512))
    whole = int(device.u32(device.crc32(t))[0])
    acc = 0
    lib = gpu_ec.lib()
    for i, cc in enumerate(chunks[:-1]):
        acc = lib.gpu_ec_crc32_combine(acc, int(cc), 512, 0)
    acc = lib.gpu_ec_crc32_combine(acc, int(chunks[-1]), S - 512 * (len(chunks) - 1), 0)
    assert acc == whole


# ------------------------------------------------------------------ batches + threads
def test_host_batches(L, orc):
    k, m, stripe_len, n = 6, 3, (1 << 20), 12
    S = erasure.shard_len(stripe_len, k)
    data = orc.fill_synthetic(stripe_len * n, 0xEC000005)
    parity = np.empty((n, m, S), np.uint8)
    crcs = np.zeros((n, k + m), np.uint32)
    full = np.zeros(n, np.uint32)
    _lib.check(L.gpu_ec_encode_batch(data.ctypes.data, stripe_len, n, k, m, parity.ctypes.data, crcs.ctypes.data,
                                     full.ctypes.data))
    shards = np.empty((n, k + m, S), np.uint8)
    for i in range(n):
        want = orc.encode(data[i * stripe_len:(i + 1) * stripe_len], k, m)
        shards[i] = want
        assert np.array_equal(parity[i], want[k:]), i
        assert crcs[i].tolist() == [zlib.crc32(s.tobytes()) for s in want]
        assert int(full[i]) == zlib.crc32(data[i * stripe_len:(i + 1) * stripe_len].tobytes())
    broken = shards.copy()
    broken[:, [1, 4, 7]] = 0
    present = np.array([1, 0, 1, 1, 0, 1, 1, 0, 1], np.uint8)
    _lib.check(L.gpu_ec_reconstruct_batch(broken.ctypes.data, present.ctypes.data, S, n, k, m))
    assert np.array_equal(broken, shards)


def test_concurrent_batch_callers(L, orc):
    """Several threads inside the _batch entry points at once (each wants 3 of a device's 4 slots): the
    slots are leased all-or-nothing, so the callers queue instead of deadlocking."""
    k, m, stripe_len, n = 6, 3, 300_000, 6
    S = erasure.shard_len(stripe_len, k)
    errs = []

    def work(tid):
        try:
            data = orc.fill_synthetic(stripe_len * n, 1000 + tid)
            parity = np.empty((n, m, S), np.uint8)
            for _ in range(3):
                _lib.check(L.gpu_ec_encode_batch(data.ctypes.data, stripe_len, n, k, m, parity.ctypes.data, None, None))
            want = orc.encode(data[:stripe_len], k, m)
            assert np.array_equal(parity[0], want[k:])
            assert checksum.crc32(data[:70_000]) == zlib.crc32(data[:70_000].tobytes())
        except Exception as e:  # noqa: BLE001
            errs.append((tid, repr(e)))

    th = [threading.Thread(target=work, args=(i,)) for i in range(6)]
    [t.start() for t in th]
    [t.join(timeout=120) for t in th]
    assert not any(t.is_alive() for t in th), "batch callers deadlocked"
    assert not errs, errs


def test_pinned_alloc_roundtrip(L, orc):
    p = C.c_void_p()
    n = 1 << 20
    _lib.check(L.gpu_ec_alloc_pinned(C.byref(p), n))
    buf = np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(n,))
    buf[:] = orc.fill_synthetic(n, 9)
    assert checksum.crc32(buf) == zlib.crc32(buf.tobytes())
    _lib.check(L.gpu_ec_free_pinned(p))


def test_concurrent_callers(L, orc):
    """Callers are arbitrary tokio worker threads: entry points must be re-entrant."""
    errs = []

    def work(tid):
        try:
            for it in range(6):
                n = 100_000 + 7919 * tid + it
                d = rnd(n, tid * 100 + it)
                want = orc.encode(d, 6, 3)
                got = erasure.encode(d, 6, 3)
                assert all(np.array_equal(a, b) for a, b in zip(got, want))
                assert checksum.crc32(d) == zlib.crc32(d.tobytes())
                opt = [None if i in (0, 5, 6) else want[i].copy() for i in range(9)]
                assert np.array_equal(erasure.decode(opt, 6, 3, n), d)
        except Exception as e:  # noqa: BLE001
            errs.append((tid, repr(e)))

    th = [threading.Thread(target=work, args=(i,)) for i in range(8)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs


# ------------------------------------------------------------------ callers either side of the path
def test_scrubber_batched_verify(L, orc, tmp_path):
    """run_background_scrubber (chunkserver.rs:642-718) as one batched verify over the storage dir."""
    store = checksum.BlockStore(str(tmp_path))
    sizes = {"a": 61, "b": 512, "c": 65_536, "d": 1_000_003, "e": 0, "f": 5 << 20, "g": 600}
    for bid, n in sizes.items():
        store.write_block(bid, rnd(n, n + 1))
    assert store.scrub() == (7, [])
    raw = bytearray((tmp_path / "d").read_bytes())
    raw[777_777] ^= 0x10
    (tmp_path / "d").write_bytes(bytes(raw))
    (tmp_path / "g").write_bytes((tmp_path / "g").read_bytes()[:-1])     # count still matches, CRC does not
    (tmp_path / "c").write_bytes((tmp_path / "c").read_bytes() + b"x")   # one more chunk than the .meta has
    (tmp_path / "b.meta").unlink()                                       # "Checksum file missing"
    assert store.scrub() == (7, ["b", "c", "d", "g"])
    blocks = [np.frombuffer((tmp_path / b).read_bytes(), dtype=np.uint8) for b in ("a", "c", "d", "f", "g")]
    metas = [(tmp_path / (b + ".meta")).read_bytes() for b in ("a", "c", "d", "f", "g")]
    got = checksum.verify_blocks(blocks, metas)
    assert got == [None, "Checksum count mismatch", f"Checksum mismatch at chunk {777_777 // 512}", None,
                   "Checksum mismatch at chunk 1"]
    # same verdicts as the block-by-block path
    assert got == [checksum.verify_chunks(b, m) for b, m in zip(blocks, metas)]


def test_write_block_checked_single_pass(L, tmp_path):
    store = checksum.BlockStore(str(tmp_path))
    data = rnd(300_000, 42)
    crc = zlib.crc32(data.tobytes())
    assert store.write_block_checked("ok", data, crc) is None
    assert store.write_block_checked("skip", data, 0) is None  # 0 = not provided (chunkserver.rs:746)
    store.verify_block("ok", data)
    assert (tmp_path / "ok.meta").read_bytes() == (tmp_path / "skip.meta").read_bytes()
    assert store.write_block_checked("bad", data, crc ^ 1) == f"Checksum mismatch: expected {crc ^ 1}, actual {crc}"
    assert not (tmp_path / "bad").exists()


def test_client_ec_write_read_orchestration(L, orc):
    from gpu_ec import client
    buf = rnd(1_000_001, 7)
    plan = client.prepare_ec_write(buf, 6, 3)
    want = orc.encode(buf, 6, 3)
    assert len(plan.shards) == 9 and plan.actual_size == buf.size
    assert all(np.array_equal(a, b) for a, b in zip(plan.shards, want))
    assert plan.shard_checksums == [zlib.crc32(s.tobytes()) for s in want]
    assert plan.full_checksum == zlib.crc32(buf.tobytes())
    # fast path: all data shards present (client/src/mod.rs:1152-1161, test :1637-1650)
    out = client.read_ec_block("blk", plan.shards[:6] + [None, None, None], 6, 3, buf.size)
    assert np.array_equal(out, buf)
    # degraded read (test :1652-1664)
    opt = [s.copy() for s in plan.shards]
    opt[0] = None
    opt[5] = None
    opt[8] = None
    assert np.array_equal(client.read_ec_block("blk", opt, 6, 3, buf.size), buf)
    with pytest.raises(erasure.ErasureError, match="EC block blk unrecoverable: only 5/6 shards available"):
        client.read_ec_block("blk", [None] * 4 + list(plan.shards[4:]), 6, 3, buf.size)


def test_two_devices_in_one_process(L, orc):
    """stripe i -> device i mod G inside the _batch entry points (no collective). Needs >= 2 GPUs."""
    if L.gpu_ec_device_count() < 2:
        pytest.skip("needs two GPUs")
    devs = (C.c_int * 2)(0, 1)
    _lib.check(L.gpu_ec_set_devices(devs, 2))
    try:
        k, m, stripe_len, n = 6, 3, 1 << 20, 10
        S = erasure.shard_len(stripe_len, k)
        data = orc.fill_synthetic(stripe_len * n, 0xEC0000D2)
        parity = np.empty((n, m, S), np.uint8)
        crcs = np.zeros((n, k + m), np.uint32)
        _lib.check(L.gpu_ec_encode_batch(data.ctypes.data, stripe_len, n, k, m, parity.ctypes.data, crcs.ctypes.data, None))
        for i in range(n):
            want = orc.encode(data[i * stripe_len:(i + 1) * stripe_len], k, m)
            assert np.array_equal(parity[i], want[k:]), i
            assert crcs[i].tolist() == [zlib.crc32(s.tobytes()) for s in want]
        # reconstruct batch across both devices (chunkserver.rs:620-623 per stripe)
        shards = np.empty((n, k + m, S), np.uint8)
        for i in range(n):
            shards[i] = orc.encode(data[i * stripe_len:(i + 1) * stripe_len], k, m)
        broken = shards.copy()
        broken[:, [1, 4, 7]] = 0
        present = np.array([1, 0, 1, 1, 0, 1, 1, 0, 1], np.uint8)
        _lib.check(L.gpu_ec_reconstruct_batch(broken.ctypes.data, present.ctypes.data, S, n, k, m))
        assert np.array_equal(broken, shards)
        # batched verify across both devices (chunkserver.rs:275-289 per block): verdicts per block
        blocks = [rnd(sz, 40 + j) for j, sz in enumerate((61, 65_536, 1_000_003, 3 << 20, 600, 2_000_000, 512))]
        metas = [orc.meta_encode_be(orc.crc32_chunks(b)) for b in blocks]
        bad = blocks[2].copy()
        bad[777_777] ^= 1
        blocks[2] = bad
        got = checksum.verify_blocks(blocks, metas)
        assert got == [None, None, f"Checksum mismatch at chunk {777_777 // 512}", None, None, None, None]
        for i in range(6):  # single calls round-robin over the devices (lease_any) too
            d = rnd(70_001 + i, i)
            assert checksum.crc32(d) == zlib.crc32(d.tobytes())
            sh = erasure.encode(d, 4, 2)
            want = orc.encode(d, 4, 2)
            assert all(np.array_equal(a, b) for a, b in zip(sh, want))
            opt = [None if j in (0, 5) else want[j].copy() for j in range(6)]
            assert np.array_equal(erasure.decode(opt, 4, 2, d.size), d)
    finally:
        one = (C.c_int * 1)(0)
        _lib.check(L.gpu_ec_set_devices(one, 1))


def test_set_devices_ignores_repeated_ordinals(L, orc):
    """{0, 0} must not make a batch lease device 0's slots twice (it would wait on itself forever)."""
    devs = (C.c_int * 2)(0, 0)
    _lib.check(L.gpu_ec_set_devices(devs, 2))
    try:
        k, m, stripe_len, n = 4, 2, 100_000, 5
        S = erasure.shard_len(stripe_len, k)
        data = orc.fill_synthetic(stripe_len * n, 0xEC0000D3)
        parity = np.empty((n, m, S), np.uint8)
        done = []
        th = threading.Thread(target=lambda: done.append(L.gpu_ec_encode_batch(data.ctypes.data, stripe_len, n, k, m, parity.ctypes.data, None, None)))
        th.start()
        th.join(timeout=60)
        assert not th.is_alive(), "batch call hung on a repeated device ordinal"
        assert done == [0]
        assert np.array_equal(parity[0], orc.encode(data[:stripe_len], k, m)[k:])
    finally:
        one = (C.c_int * 1)(0)
        _lib.check(L.gpu_ec_set_devices(one, 1))


def test_reconstruct_ec_shard_repair_path(L, orc, tmp_path):
    """chunkserver.rs:503-640 after the fetch: rebuild the target shard and store it with its .meta."""
    store = checksum.BlockStore(str(tmp_path))
    data = rnd(2_000_003, 11)
    sh = orc.encode(data, 6, 3)
    fetched = [s.copy() for s in sh]
    fetched[2] = None              # unavailable source
    fetched[7] = None
    n = store.reconstruct_ec_shard("blk_s4", 4, 6, 3, fetched)   # our own (stale) copy of shard 4 is ignored
    assert n == sh.shape[1]
    assert (tmp_path / "blk_s4").read_bytes() == sh[4].tobytes()
    store.verify_block("blk_s4", sh[4])
    assert (tmp_path / "blk_s4.meta").read_bytes() == orc.meta_encode_be(orc.crc32_chunks(sh[4]))
    with pytest.raises(ValueError, match="ec_shard_sources length 8 != total shards 9"):
        store.reconstruct_ec_shard("x", 0, 6, 3, fetched[:8])
    fetched[0] = fetched[1] = None  # now only 3, 5, 6, 8 present besides the target
    with pytest.raises(ValueError, match="Only 4 shards available, need at least 6 for reconstruction"):
        store.reconstruct_ec_shard("x", 4, 6, 3, fetched)


def test_reconstruct_only_wanted_slots(L, orc):
    data = rnd(500_003, 3)
    for k, m, missing, want in [(6, 3, [1, 4, 7], [4]), (6, 3, [0, 8], [8, 0]), (10, 4, [2, 5, 11, 13], [13, 2]), (4, 2, [3], [3])]:
        sh = orc.encode(data, k, m)
        opt = [None if i in missing else sh[i].copy() for i in range(k + m)]
        erasure.reconstruct_only(opt, k, m, want)
        for i in range(k + m):
            if i in want:
                assert np.array_equal(opt[i], sh[i]), (k, m, i)
            elif i in missing:
                assert opt[i] is None
    opt = [s.copy() for s in orc.encode(data, 6, 3)]
    opt[2] = None
    with pytest.raises(erasure.ErasureError, match="invalid argument"):
        erasure.reconstruct_only(opt, 6, 3, [3])  # slot 3 is present


def test_convert_stored_block_to_ec(L, orc, tmp_path):
    """SURVEY §8(f) item 3 (master.rs:2108-2113 TODO): replicated block -> RS(6,3) shards on the chunkserver."""
    store = checksum.BlockStore(str(tmp_path))
    data = rnd(3_000_001, 21)
    store.write_block("cold", data)
    info = store.convert_block_to_ec("cold", 6, 3)
    want = orc.encode(data, 6, 3)
    assert info["original_size"] == data.size and info["full_checksum"] == zlib.crc32(data.tobytes())
    assert info["shard_checksums"] == [zlib.crc32(s.tobytes()) for s in want]
    for i, sid in enumerate(info["shard_ids"]):
        raw = np.frombuffer((tmp_path / sid).read_bytes(), dtype=np.uint8)
        assert np.array_equal(raw, want[i])
        store.verify_block(sid, raw)  # the .meta written from the encode call verifies like any block
    assert store.scrub() == (10, [])
    # degraded read from the stored shards (client/src/mod.rs:1110-1165)
    from gpu_ec import client
    opt = [None if i in (0, 4, 8) else np.frombuffer((tmp_path / sid).read_bytes(), dtype=np.uint8)
           for i, sid in enumerate(info["shard_ids"])]
    assert np.array_equal(client.read_ec_block("cold", opt, 6, 3, info["original_size"]), data)
