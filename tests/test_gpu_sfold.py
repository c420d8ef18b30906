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
    os.environ["GPU_EC_SFOLD"] = "1" if sfold else "0"
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
