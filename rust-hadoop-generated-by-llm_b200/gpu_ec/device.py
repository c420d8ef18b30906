"""Device-resident batches for bench.py and the GPU tests.

torch is used ONLY for device memory, streams and events (plumbing); every byte of compute goes
through the `_dev` entry points of libgpu_ec.so. Stripes live in one pitched uint8 tensor
[n_stripes, k+m, pitch] (pitch = shard_len rounded up to 256) — the HBM layout DESIGN.md describes.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from ._lib import POLY_IEEE, Stripes, check, lib


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def pitch_of(shard_len: int) -> int:
    return (shard_len + 255) // 256 * 256


class StripeBatch:
    """n_stripes x (k+m) shard slots in HBM."""

    def __init__(self, n_stripes: int, k: int, m: int, shard_len: int, device="cuda"):
        self.n_stripes, self.k, self.m, self.shard_len = n_stripes, k, m, shard_len
        self.pitch = pitch_of(shard_len)
        # uninitialised on purpose: fill_synthetic() writes every byte, encode() every parity byte
        self.buf = torch.empty((n_stripes, k + m, self.pitch), dtype=torch.uint8, device=device)

    def desc(self, first: int = 0, count: int | None = None) -> Stripes:
        count = self.n_stripes - first if count is None else count
        stripe_stride = (self.k + self.m) * self.pitch
        return Stripes(self.buf.data_ptr() + first * stripe_stride, self.pitch, stripe_stride, self.shard_len, count)

    def fill_synthetic(self, seed: int, original_len: int = 0) -> None:
        """One launch fills the whole pitched buffer with the synthetic stream (slot s of stripe i
        starts at stream position (i*(k+m) + s) * pitch, so the oracle can regenerate any shard
        independently); parity slots are overwritten by encode. With original_len < k*shard_len
        the tail of the data slots is zeroed like erasure.rs:18 pads it."""
        check(lib().gpu_ec_fill_synthetic_dev(self.buf.data_ptr(), self.buf.numel(), seed, 0, _stream_ptr()))
        if original_len and original_len < self.k * self.shard_len:
            for j in range(self.k):
                keep = min(max(original_len - j * self.shard_len, 0), self.shard_len)
                if keep < self.shard_len:
                    self.buf[:, j, keep:self.shard_len] = 0

    def stream_offset(self, i: int, s: int) -> int:
        return (i * (self.k + self.m) + s) * self.pitch

    def shard_host(self, i: int, s: int) -> np.ndarray:
        return self.buf[i, s, :self.shard_len].cpu().numpy()

    def encode(self) -> None:
        d = self.desc()
        check(lib().gpu_ec_encode_dev(C.byref(d), self.k, self.m, _stream_ptr()))

    def reconstruct(self, present) -> None:
        pres = np.ascontiguousarray(present, dtype=np.uint8)
        d = self.desc()
        check(lib().gpu_ec_reconstruct_dev(C.byref(d), pres.ctypes.data, self.k, self.m, _stream_ptr()))

    def encode_crc(self, original_len: int = 0, poly: int = POLY_IEEE, want_full: bool = True):
        n = self.k + self.m
        shard_crc = torch.empty((self.n_stripes, n), dtype=torch.int32, device=self.buf.device)
        full = torch.empty(self.n_stripes, dtype=torch.int32, device=self.buf.device) if want_full else None
        d = self.desc()
        check(lib().gpu_ec_encode_crc_dev(C.byref(d), self.k, self.m, original_len, poly, shard_crc.data_ptr(),
                                          full.data_ptr() if want_full else None, _stream_ptr()))
        return shard_crc, full

    def shard_crcs(self, poly: int = POLY_IEEE) -> torch.Tensor:
        n = self.k + self.m
        out = torch.empty((self.n_stripes, n), dtype=torch.int32, device=self.buf.device)
        check(lib().gpu_ec_crc32_dev(self.buf.data_ptr(), self.n_stripes * n, self.pitch, self.shard_len, poly,
                                     out.data_ptr(), _stream_ptr()))
        return out


def fill_synthetic(t: torch.Tensor, seed: int, byte_offset: int = 0) -> None:
    check(lib().gpu_ec_fill_synthetic_dev(t.data_ptr(), t.numel(), seed, byte_offset, _stream_ptr()))


def crc32_chunks(t: torch.Tensor, chunk: int, poly: int = POLY_IEEE, out: torch.Tensor | None = None,
                 big_endian: bool = False) -> torch.Tensor:
    """t: 1-D uint8 device tensor -> int32 tensor of ceil(len/chunk) CRCs (bit pattern = u32)."""
    n = (t.numel() + chunk - 1) // chunk
    if out is None:
        out = torch.empty(n, dtype=torch.int32, device=t.device)
    check(lib().gpu_ec_crc32_chunks_dev(t.data_ptr(), 1, (t.numel() + 15) // 16 * 16, t.numel(), chunk, poly,
                                        out.data_ptr(), n, int(big_endian), _stream_ptr()))
    return out


def crc32(t: torch.Tensor, poly: int = POLY_IEEE, out: torch.Tensor | None = None) -> torch.Tensor:
    if out is None:
        out = torch.empty(1, dtype=torch.int32, device=t.device)
    check(lib().gpu_ec_crc32_dev(t.data_ptr(), 1, (t.numel() + 15) // 16 * 16, t.numel(), poly, out.data_ptr(), _stream_ptr()))
    return out


def verify_chunks(t: torch.Tensor, chunk: int, meta_be: torch.Tensor, poly: int = POLY_IEEE,
                  out: torch.Tensor | None = None) -> torch.Tensor:
    """-> int64 device tensor [1]: first bad chunk or -1."""
    if out is None:
        out = torch.empty(1, dtype=torch.int64, device=t.device)
    check(lib().gpu_ec_verify_chunks_dev(t.data_ptr(), t.numel(), chunk, poly, meta_be.data_ptr(), out.data_ptr(), _stream_ptr()))
    return out


def u32(t: torch.Tensor) -> np.ndarray:
    return t.cpu().numpy().view(np.uint32)
