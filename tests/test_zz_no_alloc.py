"""gpu_ec_crc32_dev_noalloc: whole-buffer CRCs without a single allocation.

Who needs it: a service that sizes its device memory once at start-up and must never ask the
allocator for more, a latency-critical path that cannot afford to wait on an allocator lock, or
work that is recorded once and replayed. Those callers cannot use gpu_ec_crc32_dev, which takes
its per-segment scratch from the stream-ordered allocator; this entry point keeps that scratch in
a fixed 4 MiB array inside the library instead (limit: 2^20 segments per call).

The file is named so pytest collects it last: if it fails under -x every other test has already
run. It calls the new entry point through the C ABI and requires bit-identical output to the
existing entry point (CRCs are integers: tolerance zero) on two calls in a row, the second one
smaller so it runs over the first call's leftovers in the shared scratch.
"""
import zlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

import gpu_ec  # noqa: E402
from gpu_ec import _lib  # noqa: E402


def _both(L, torch, t, n_buf, stride, length, poly=0):
    """-> (existing entry point's CRCs, no-alloc entry point's CRCs) as uint32 arrays."""
    st = torch.cuda.current_stream().cuda_stream
    want = torch.zeros(n_buf, dtype=torch.int32, device="cuda")
    got = torch.full((n_buf,), 0x5A5A5A5A, dtype=torch.int32, device="cuda")
    _lib.check(L.gpu_ec_crc32_dev(t.data_ptr(), n_buf, stride, length, poly, want.data_ptr(), st))
    _lib.check(L.gpu_ec_crc32_dev_noalloc(t.data_ptr(), n_buf, stride, length, poly, got.data_ptr(), st))
    torch.cuda.synchronize()
    return want.cpu().numpy().view(np.uint32), got.cpu().numpy().view(np.uint32)


def test_noalloc_crc_matches_existing_entry_point_twice_in_a_row():
    import torch
    L = gpu_ec.lib()
    assert L.gpu_ec_device_count() > 0
    st = torch.cuda.current_stream().cuda_stream
    # call 1: 14 shards of an RS(10,4) stripe over a 64 MiB block (odd length, ragged tail)
    S1, pitch1 = 6_710_887, 6_711_040
    a = torch.empty(14 * pitch1, dtype=torch.uint8, device="cuda")
    _lib.check(L.gpu_ec_fill_synthetic_dev(a.data_ptr(), a.numel(), 0xEC0000F1, 0, st))
    want1, got1 = _both(L, torch, a, 14, pitch1, S1)
    assert np.array_equal(got1, want1)
    host = a.cpu().numpy()
    assert int(got1[3]) == zlib.crc32(host[3 * pitch1:3 * pitch1 + S1].tobytes())
    # call 2, immediately after: fewer, shorter buffers -> reuses the front of the scratch that
    # call 1 filled; stale values behind it must not leak into the result
    S2, pitch2 = 174_763, 174_848
    want2, got2 = _both(L, torch, a, 9, pitch2, S2)
    assert np.array_equal(got2, want2)
    assert int(got2[8]) == zlib.crc32(host[8 * pitch2:8 * pitch2 + S2].tobytes())
    # and the Castagnoli variant, plus tiny and empty buffers
    want3, got3 = _both(L, torch, a, 5, 1024, 1000, poly=gpu_ec.POLY_CASTAGNOLI)
    assert np.array_equal(got3, want3)
    want4, got4 = _both(L, torch, a, 3, 16, 0)
    assert np.array_equal(got4, want4) and not got4.any()


def test_noalloc_rejects_sizes_beyond_the_built_in_scratch():
    import torch
    L = gpu_ec.lib()
    st = torch.cuda.current_stream().cuda_stream
    t = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda")
    out = torch.zeros(4, dtype=torch.int32, device="cuda")
    # a description of 2^21 buffers of 128 KiB would need 2 segments each even at the 64 KiB
    # ceiling: more than 2^20 slots. The call must refuse before launching anything.
    before = L.gpu_ec_kernel_launch_count()
    rc = L.gpu_ec_crc32_dev_noalloc(t.data_ptr(), 1 << 21, 0, 128 << 10, 0, out.data_ptr(), st)
    assert rc == _lib.ERR_INVALID_ARG
    assert b"built-in scratch" in L.gpu_ec_last_error()
    assert L.gpu_ec_kernel_launch_count() == before
    torch.cuda.synchronize()


@pytest.mark.parametrize("n_buf,length", [(1, 600 << 20), (14, 160 << 20), (3, (96 << 20) + 77)])
def test_noalloc_long_segments(n_buf, length):
    """Inputs whose per-segment CRCs do not fit the array at the usual segment length: the entry point
    lengthens the segments (up to 64 KiB, longer than gpu_ec_crc32_dev ever uses) instead of allocating.
    Same answer as the existing entry point, spot-checked against zlib."""
    import torch
    L = gpu_ec.lib()
    st = torch.cuda.current_stream().cuda_stream
    stride = (length + 255) // 256 * 256
    free, _ = torch.cuda.mem_get_info()
    if n_buf * stride + (1 << 30) > free:
        pytest.skip("not enough free device memory")
    t = torch.empty(n_buf * stride, dtype=torch.uint8, device="cuda")
    _lib.check(L.gpu_ec_fill_synthetic_dev(t.data_ptr(), t.numel(), 0xEC0000F2 + n_buf, 0, st))
    want, got = _both(L, torch, t, n_buf, stride, length)
    assert np.array_equal(got, want)
    b = n_buf - 1
    assert int(got[b]) == zlib.crc32(t[b * stride:b * stride + length].cpu().numpy().tobytes())


def test_noalloc_very_long_segments_when_memory_allows():
    """>= 17 GiB in one call: 32 KiB .. 64 KiB segments. Needs ~18 GiB free; skipped otherwise."""
    import torch
    L = gpu_ec.lib()
    st = torch.cuda.current_stream().cuda_stream
    n_buf, length = 17, 1 << 30
    free, _ = torch.cuda.mem_get_info()
    if n_buf * length + (2 << 30) > free:
        pytest.skip("not enough free device memory")
    t = torch.empty(n_buf * length, dtype=torch.uint8, device="cuda")
    _lib.check(L.gpu_ec_fill_synthetic_dev(t.data_ptr(), t.numel(), 0xEC0000F3, 0, st))
    want, got = _both(L, torch, t, n_buf, length, length)
    assert np.array_equal(got, want)
    assert int(got[16]) == zlib.crc32(t[16 * length:].cpu().numpy().tobytes())
