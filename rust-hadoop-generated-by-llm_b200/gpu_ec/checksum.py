"""Host-side mirror of the chunkserver's checksum functions
(dfs/chunkserver/src/chunkserver.rs:16,182-196,238-351,746-766), same names and error strings.

The reference keeps `<block_id>` and `<block_id>.meta` (big-endian u32 per 512-byte chunk) in a
storage directory; `BlockStore` below reproduces that layout so verify_block / verify_partial_read
read like the originals. CRCs are computed by libgpu_ec.so's CUDA kernels.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import numpy as np

from ._lib import ERR_COUNT_MISMATCH, ERR_CRC_MISMATCH, POLY_IEEE, check, lib

CHECKSUM_CHUNK_SIZE = 512  # chunkserver.rs:16


def _as_u8(data) -> np.ndarray:
    if isinstance(data, np.ndarray):
        return np.ascontiguousarray(data, dtype=np.uint8).reshape(-1)
    return np.frombuffer(bytes(data), dtype=np.uint8)


def crc32(data, poly: int = POLY_IEEE) -> int:
    """crc32fast::Hasher::new(); update(data); finalize()  (chunkserver.rs:746-750)."""
    d = _as_u8(data)
    out = C.c_uint32(0)
    check(lib().gpu_ec_crc32(d.ctypes.data if d.size else None, d.size, poly, C.byref(out)))
    return out.value


def calculate_checksums(data, chunk: int = CHECKSUM_CHUNK_SIZE, poly: int = POLY_IEEE) -> np.ndarray:
    """chunkserver.rs:182-190 -> one CRC-32 per 512-byte chunk (last chunk short; empty -> empty)."""
    d = _as_u8(data)
    n = (d.size + chunk - 1) // chunk
    out = np.empty(n, dtype=np.uint32)
    check(lib().gpu_ec_crc32_chunks(d.ctypes.data if d.size else None, d.size, chunk, poly, out.ctypes.data if n else None))
    return out


def meta_bytes(checksums) -> bytes:
    """chunkserver.rs:196: checksums.iter().flat_map(|c| c.to_be_bytes())."""
    c = np.ascontiguousarray(checksums, dtype=np.uint32)
    out = np.empty(4 * c.size, dtype=np.uint8)
    lib().gpu_ec_meta_encode_be(c.ctypes.data, c.size, out.ctypes.data)
    return out.tobytes()


def block_checksums(data, chunk: int = CHECKSUM_CHUNK_SIZE, poly: int = POLY_IEEE):
    """Whole-block CRC (chunkserver.rs:746-750) and per-chunk CRCs (:195) from one pass."""
    d = _as_u8(data)
    n = (d.size + chunk - 1) // chunk
    out = np.empty(n, dtype=np.uint32)
    whole = C.c_uint32(0)
    check(lib().gpu_ec_block_checksums(d.ctypes.data if d.size else None, d.size, chunk, poly, C.byref(whole),
                                       out.ctypes.data if n else None))
    return whole.value, out


def verify_chunks(data, meta_be: bytes, chunk: int = CHECKSUM_CHUNK_SIZE, poly: int = POLY_IEEE) -> Optional[str]:
    """The compare of chunkserver.rs:265-289 -> None when OK, else the reference's error string."""
    d = _as_u8(data)
    mb = np.frombuffer(bytes(meta_be), dtype=np.uint8)
    bad = C.c_int64(-1)
    rc = lib().gpu_ec_verify_chunks(d.ctypes.data if d.size else None, d.size, chunk, poly,
                                    mb.ctypes.data if mb.size else None, mb.size, C.byref(bad))
    if rc == ERR_COUNT_MISMATCH:
        return "Checksum count mismatch"
    if rc == ERR_CRC_MISMATCH:
        return f"Checksum mismatch at chunk {bad.value}"
    check(rc)
    return None


def check_expected(data, expected_checksum: int) -> Optional[str]:
    """WriteBlock's in-flight check (chunkserver.rs:746-766): 0 means "not provided" -> skip."""
    if expected_checksum == 0:
        return None
    actual = crc32(data)
    if actual != expected_checksum:
        return f"Checksum mismatch: expected {expected_checksum}, actual {actual}"
    return None


class BlockStore:
    """The slice of MyChunkServer that owns `<dir>/<block_id>` + `<dir>/<block_id>.meta`."""

    def __init__(self, storage_dir: str, cold_storage_dir: Optional[str] = None):
        self.storage_dir = storage_dir
        self.cold_storage_dir = cold_storage_dir

    def write_block(self, block_id: str, data) -> None:
        """write_block_async (chunkserver.rs:192-209): data file + big-endian .meta sidecar."""
        d = _as_u8(data)
        meta = meta_bytes(calculate_checksums(d))
        with open(os.path.join(self.storage_dir, block_id), "wb") as f:
            f.write(d.tobytes())
        with open(os.path.join(self.storage_dir, block_id + ".meta"), "wb") as f:
            f.write(meta)

    def _meta_path(self, block_id: str) -> str:
        hot = os.path.join(self.storage_dir, block_id + ".meta")
        if os.path.exists(hot):
            return hot
        if self.cold_storage_dir:
            cold = os.path.join(self.cold_storage_dir, block_id + ".meta")
            if os.path.exists(cold):
                return cold
        return hot

    def verify_block(self, block_id: str, data) -> None:
        """chunkserver.rs:238-292. Raises ValueError(<reference message>) on failure."""
        path = self._meta_path(block_id)
        if not os.path.exists(path):
            raise ValueError("Checksum file missing")
        with open(path, "rb") as f:
            meta = f.read()
        err = verify_chunks(data, meta)
        if err:
            raise ValueError(err)

    def verify_partial_read(self, block_id: str, offset: int, length: int) -> None:
        """chunkserver.rs:296-351: only the chunks touched by [offset, offset+length), hot dir only."""
        meta_path = os.path.join(self.storage_dir, block_id + ".meta")
        if not os.path.exists(meta_path):
            raise ValueError("Checksum file missing")
        with open(meta_path, "rb") as f:
            meta = f.read()
        n_all = len(meta) // 4
        start = offset // CHECKSUM_CHUNK_SIZE
        end = max(offset + length - 1, 0) // CHECKSUM_CHUNK_SIZE
        end = min(end, n_all - 1)
        if end < start:
            return
        path = os.path.join(self.storage_dir, block_id)
        size = os.path.getsize(path)
        lo = start * CHECKSUM_CHUNK_SIZE
        hi = min((end + 1) * CHECKSUM_CHUNK_SIZE, size)
        with open(path, "rb") as f:
            f.seek(lo)
            buf = f.read(hi - lo)
        got = calculate_checksums(np.frombuffer(buf, dtype=np.uint8))
        want = np.frombuffer(meta[4 * start:4 * (start + got.size)], dtype=">u4")
        bad = np.nonzero(got != want.astype(np.uint32))[0]
        if bad.size:
            raise ValueError(f"Checksum mismatch at chunk {start + int(bad[0])}")
