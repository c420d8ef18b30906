"""GPU parity of the sparse-fold whole-buffer CRC (csrc/sfold_kernels.cuh) against zlib and against the table
kernel, through the C ABI (gpu_ec_crc32_dev). Replaces the inline whole-buffer crc32fast::Hasher sites
(dfs/chunkserver/src/chunkserver.rs:746-750; dfs/client/src/mod.rs:601-611). Bit-exact or fail."""
import os
import zlib

import numpy as np
import pytest

import gpu_ec
from gpu_ec import _lib

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def L():
    lib = gpu_ec.lib()
    if lib.gpu_ec_device_count() == 0:
        pytest.skip("no CUDA device")
    _lib.check(lib.gpu_ec_init())
    return lib


def _crc_dev(L, t, n_buf, stride, length, sfold):
    import torch
    os.environ["GPU_EC_SFOLD"] = "2" if sfold else "0"
    out = torch.zeros(n_buf, dtype=torch.int32, device="cuda")
    _lib.check(L.gpu_ec_crc32_dev(t.data_ptr(), n_buf, stride, length, 0, out.data_ptr(), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    os.environ.pop("GPU_EC_SFOLD", None)
    return out.cpu().numpy().view(np.uint32)


# (n_buf, len): every stride of the table gets picked somewhere; ragged tails (len % 16 = 1, 7, 15), lengths that
# leave whole virtual rows in front of segment 0, one buffer split into row segments, many short buffers
CASES = [(1, 8 << 20), (1, (8 << 20) + 1), (3, 6_710_887), (14, 6_710_887), (2, 11_184_811), (1, (64 << 20) + 7),
         (40, 262_144 + 15), (9, 1_000_003), (1, 300 << 20), (64, 300_001), (5, 2_097_152 - 16)]


@pytest.mark.parametrize("n_buf,length", CASES)
def test_sfold_crc_vs_zlib_and_table_kernel(L, n_buf, length):
    import torch
    from gpu_ec import device
    stride = (length + 255) // 256 * 256
    t = torch.empty(n_buf * stride, dtype=torch.uint8, device="cuda")
    device.fill_synthetic(t, 0xEC005F01 + n_buf)  # garbage (not zeros) between len and the stride: the tail mask matters
    got = _crc_dev(L, t, n_buf, stride, length, sfold=True)
    ref = _crc_dev(L, t, n_buf, stride, length, sfold=False)
    assert np.array_equal(got, ref)
    host = t.cpu().numpy().reshape(n_buf, stride)
    for b in sorted({0, n_buf // 2, n_buf - 1}):
        assert int(got[b]) == zlib.crc32(host[b, :length].tobytes()), b


def test_sfold_adversarial_fills(L):
    import torch
    n_buf, length = 4, 3_000_001
    stride = (length + 255) // 256 * 256
    for fill in (0x00, 0xFF):
        t = torch.full((n_buf * stride,), fill, dtype=torch.uint8, device="cuda")
        got = _crc_dev(L, t, n_buf, stride, length, sfold=True)
        assert all(int(c) == zlib.crc32(bytes([fill]) * length) for c in got)
    t = (torch.arange(n_buf * stride, device="cuda") % 256).to(torch.uint8)  # the reference's own pattern (chunkserver.rs:1142-1144)
    got = _crc_dev(L, t, n_buf, stride, length, sfold=True)
    host = t.cpu().numpy().reshape(n_buf, stride)
    assert [int(c) for c in got] == [zlib.crc32(host[b, :length].tobytes()) for b in range(n_buf)]


# ---------------------------------------------------------------- fused encode + shard CRCs (fused_kernels.cuh)
FUSED = [(10, 4, (8 << 20) + 5, 20), (10, 4, 64 << 20, 3), (8, 3, (4 << 20) + 1, 40), (12, 4, (6 << 20) - 7, 24),
         (10, 2, (8 << 20) + 16, 16), (9, 4, 5_000_011, 33), (11, 1, 9_999_999, 17), (8, 4, (4 << 20) + 33, 30),
         (10, 4, 3_000_000, 75), (6, 3, (64 << 20) + 1, 6), (6, 3, 4_000_000, 160), (4, 2, 2_000_003, 200), (7, 4, 8_000_000, 40),
         (2, 2, 1_500_000, 300), (5, 1, 3_333_333, 150), (1, 1, 1_000_000, 400), (3, 3, (8 << 20) + 9, 50)]


@pytest.mark.parametrize("k,m,stripe_len,n_stripes", FUSED)
def test_fused_encode_crc_vs_oracle_and_two_pass(L, k, m, stripe_len, n_stripes):
    """client write path dfs/client/src/mod.rs:600-611 in one pass: parity, k+m shard CRCs and CRC(original)
    must equal the oracle's and the two-pass path's, bit for bit."""
    import torch
    from gpu_ec import device, erasure
    from oracle import oracle as orc
    S = erasure.shard_len(stripe_len, k)
    b = device.StripeBatch(n_stripes, k, m, S)
    b.fill_synthetic(0xEC00F05D + k, original_len=stripe_len)
    os.environ["GPU_EC_SFOLD"] = "0"
    crc_ref, full_ref = b.encode_crc(original_len=stripe_len)
    torch.cuda.synchronize()
    par_ref = b.buf[:, k:, :S].clone()
    b.buf[:, k:] = 0x5A
    os.environ["GPU_EC_SFOLD"] = "1"
    l0 = L.gpu_ec_kernel_launch_count()
    crc, full = b.encode_crc(original_len=stripe_len)
    torch.cuda.synchronize()
    os.environ.pop("GPU_EC_SFOLD", None)
    assert L.gpu_ec_kernel_launch_count() - l0 <= 6, "expected the single-pass path (fused kernel + finisher launches)"
    assert torch.equal(b.buf[:, k:, :S], par_ref)
    assert np.array_equal(device.u32(crc), device.u32(crc_ref))
    assert np.array_equal(device.u32(full), device.u32(full_ref))
    crc = device.u32(crc)
    for i in sorted({0, n_stripes - 1}):
        data = np.concatenate([b.shard_host(i, j) for j in range(k)])
        want = orc.encode(data, k, m)
        for s in range(k, k + m):
            assert np.array_equal(b.shard_host(i, s), want[s]), (i, s)
        assert crc[i].tolist() == [zlib.crc32(x.tobytes()) for x in want]
        assert int(device.u32(full)[i]) == zlib.crc32(data[:stripe_len].tobytes())


@pytest.mark.parametrize("k,m,n", [(10, 4, (64 << 20) + 3), (6, 3, 20_000_003), (10, 4, 1_000_001)])
def test_host_encode_with_checksums_all_paths(L, k, m, n):
    """gpu_ec_encode through host pointers: (a) shard CRCs only -> the fused pass where the shape allows it,
    (b) with .meta arrays -> whole-shard CRCs folded from the per-512-byte CRCs (no extra pass). Both against
    the oracle (dfs/client/src/mod.rs:600-611; dfs/chunkserver/src/chunkserver.rs:182-196)."""
    from gpu_ec import erasure
    from oracle import oracle as orc
    data = orc.fill_synthetic(n, 0xEC00A0A0 + k)
    want = orc.encode(data, k, m)
    want_crc = [zlib.crc32(s.tobytes()) for s in want]
    for with_meta in (False, True):
        shards, crcs, full, metas = erasure.encode_with_checksums(data, k, m, with_meta=with_meta)
        for i in range(k + m):
            assert np.array_equal(shards[i], want[i]), (with_meta, i)
        assert crcs.tolist() == want_crc, with_meta
        assert full == zlib.crc32(data.tobytes())
        if with_meta:
            for i in (0, k - 1, k, k + m - 1):
                assert metas[i].tobytes() == orc.meta_encode_be(orc.crc32_chunks(want[i]))


def test_fused_with_gaps_between_stripes_and_shards(L):
    """gpu_ec_stripes with a shard_stride larger than the pitch and a stripe_stride larger than (k+m) shards:
    the fused pass must address through the strides it is given, not assume a packed batch."""
    import ctypes as C
    import torch
    from gpu_ec import device, erasure
    from gpu_ec._lib import Stripes
    from oracle import oracle as orc
    k, m, stripe_len, n = 10, 4, (4 << 20) + 7, 40
    S = erasure.shard_len(stripe_len, k)
    shard_stride = (S + 255) // 256 * 256 + 512
    stripe_stride = (k + m + 1) * shard_stride + 4096
    buf = torch.empty(n * stripe_stride, dtype=torch.uint8, device="cuda")
    device.fill_synthetic(buf, 0xEC0000C7)
    view = buf.view(n, stripe_stride)
    pad = S * k - stripe_len
    if pad:
        view[:, (k - 1) * shard_stride + S - pad:(k - 1) * shard_stride + S] = 0
    d = Stripes(buf.data_ptr(), shard_stride, stripe_stride, S, n)
    out = {}
    for mode in ("0", "1"):
        os.environ["GPU_EC_SFOLD"] = mode
        crc = torch.zeros((n, k + m), dtype=torch.int32, device="cuda")
        full = torch.zeros(n, dtype=torch.int32, device="cuda")
        _lib.check(L.gpu_ec_encode_crc_dev(C.byref(d), k, m, stripe_len, 0, crc.data_ptr(), full.data_ptr(), torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        out[mode] = (device.u32(crc).copy(), device.u32(full).copy())
    os.environ.pop("GPU_EC_SFOLD", None)
    assert np.array_equal(out["0"][0], out["1"][0]) and np.array_equal(out["0"][1], out["1"][1])
    i = n - 1
    host = view[i].cpu().numpy()
    data = np.concatenate([host[j * shard_stride:j * shard_stride + S] for j in range(k)])[:stripe_len]
    want = orc.encode(data, k, m)
    for s in range(k, k + m):
        assert np.array_equal(host[s * shard_stride:s * shard_stride + S], want[s]), s
    assert out["1"][0][i].tolist() == [zlib.crc32(x.tobytes()) for x in want]
    assert int(out["1"][1][i]) == zlib.crc32(data.tobytes())
