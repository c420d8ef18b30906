"""Host-side mirror of the chunkserver's checksum functions
(dfs/chunkserver/src/chunkserver.rs:16,182-196,238-351,746-766), same names and error strings.

The reference keeps `<block_id>` and `<block_id>.meta` (big-endian u32 per 512-byte chunk) in a
storage directory; `BlockStore` below reproduces that layout so verify_block / verify_partial_read
read like the originals. CRCs are computed by libgpu_ec.so's CUDA kernels.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional, Sequence, Tuple

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


def block_checksums_batch(blocks: Sequence, chunk: int = CHECKSUM_CHUNK_SIZE, poly: int = POLY_IEEE):
    """block_checksums for many (small) blocks with one launch and one synchronisation: what a chunkserver does
    per WriteBlock (chunkserver.rs:746-769 + :195), batched. -> (whole CRCs uint32[n], [chunk CRC arrays])."""
    arrs = [_as_u8(b) for b in blocks]
    n = len(arrs)
    outs = [np.empty((a.size + chunk - 1) // chunk, dtype=np.uint32) for a in arrs]
    whole = np.zeros(n, dtype=np.uint32)
    dptr = (C.c_void_p * n)(*[a.ctypes.data if a.size else None for a in arrs])
    optr = (C.c_void_p * n)(*[o.ctypes.data if o.size else None for o in outs])
    lens = (C.c_size_t * n)(*[a.size for a in arrs])
    check(lib().gpu_ec_block_checksums_batch(n, dptr, lens, chunk, poly, whole.ctypes.data if n else None, optr))
    return whole, outs


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


def verify_blocks(blocks: Sequence, metas: Sequence[bytes], chunk: int = CHECKSUM_CHUNK_SIZE,
                  poly: int = POLY_IEEE) -> List[Optional[str]]:
    """Batched verify_block (the scrubber's inner loop, chunkserver.rs:664-690): one call, blocks
    pipelined over streams. -> per block None or the reference's error string."""
    n = len(blocks)
    arrs = [_as_u8(b) for b in blocks]
    mbs = [np.frombuffer(bytes(m), dtype=np.uint8) for m in metas]
    dptr = (C.c_void_p * n)(*[a.ctypes.data if a.size else None for a in arrs])
    mptr = (C.c_void_p * n)(*[m.ctypes.data if m.size else None for m in mbs])
    lens = (C.c_size_t * n)(*[a.size for a in arrs])
    mlens = (C.c_size_t * n)(*[m.size for m in mbs])
    status = (C.c_int * n)()
    bad = (C.c_int64 * n)()
    check(lib().gpu_ec_verify_chunks_batch(n, dptr, lens, chunk, poly, mptr, mlens, status, bad))
    out: List[Optional[str]] = []
    for i in range(n):
        if status[i] == ERR_COUNT_MISMATCH:
            out.append("Checksum count mismatch")
        elif status[i] == ERR_CRC_MISMATCH:
            out.append(f"Checksum mismatch at chunk {bad[i]}")
        else:
            check(status[i])
            out.append(None)
    return out


class BlockStore:
    """The slice of MyChunkServer that owns `<dir>/<block_id>` + `<dir>/<block_id>.meta`."""

    def __init__(self, storage_dir: str, cold_storage_dir: Optional[str] = None):
        self.storage_dir = storage_dir
        self.cold_storage_dir = cold_storage_dir

    def write_block_checked(self, block_id: str, data, expected_checksum: int = 0) -> Optional[str]:
        """WriteBlock's checksum work in ONE pass over the bytes: the reference
        hashes req.data twice, whole-block verify (chunkserver.rs:746-766) then per-512-byte
        (:195). -> None on success or the reference's "Checksum mismatch: ..." text (nothing written)."""
        d = _as_u8(data)
        whole, chunks = block_checksums(d)
        if expected_checksum != 0 and whole != expected_checksum:
            return f"Checksum mismatch: expected {expected_checksum}, actual {whole}"
        with open(os.path.join(self.storage_dir, block_id), "wb") as f:
            f.write(d.tobytes())
        with open(os.path.join(self.storage_dir, block_id + ".meta"), "wb") as f:
            f.write(meta_bytes(chunks))
        return None

    def scrub(self) -> Tuple[int, List[str]]:
        """run_background_scrubber's pass (chunkserver.rs:652-701): every non-.meta file of the
        storage dir is re-verified; -> (blocks scanned, ids of corrupted blocks). One batched call."""
        ids = sorted(n for n in os.listdir(self.storage_dir)
                     if not n.endswith(".meta") and os.path.isfile(os.path.join(self.storage_dir, n)))
        blocks, metas, have = [], [], []
        bad: List[str] = []
        for bid in ids:
            mp = self._meta_path(bid)
            if not os.path.exists(mp):
                bad.append(bid)  # "Checksum file missing" counts as a failed verify
                continue
            with open(os.path.join(self.storage_dir, bid), "rb") as f:
                blocks.append(np.frombuffer(f.read(), dtype=np.uint8))
            with open(mp, "rb") as f:
                metas.append(f.read())
            have.append(bid)
        for bid, err in zip(have, verify_blocks(blocks, metas)):
            if err:
                bad.append(bid)
        return len(ids), sorted(bad)

    def reconstruct_ec_shard(self, block_id: str, shard_index: int, data_shards: int, parity_shards: int,
                             fetched: Sequence) -> int:
        """The byte work of MyChunkServer::reconstruct_ec_shard (chunkserver.rs:503-640) once the peers'
        shards have been fetched: `fetched[i]` is shard i or None (unavailable source). The target slot is
        forced to None (:529), at least k others must be present (:607-617), every missing shard is
        rebuilt (:620-623) and shard `shard_index` is written with its .meta (:626-631). -> bytes written."""
        from . import erasure
        total = data_shards + parity_shards
        if len(fetched) != total:
            raise ValueError(f"ec_shard_sources length {len(fetched)} != total shards {total}")
        opt = [None if (i == shard_index or f is None) else _as_u8(f) for i, f in enumerate(fetched)]
        available = sum(o is not None for o in opt)
        if available < data_shards:
            raise ValueError(f"Only {available} shards available, need at least {data_shards} for reconstruction")
        # the reference rebuilds every missing shard and keeps one; only the kept one is computed here
        erasure.reconstruct_only(opt, data_shards, parity_shards, [shard_index])
        shard = opt[shard_index]
        if shard is None:
            raise ValueError(f"Shard {shard_index} still None after reconstruct")
        self.write_block(block_id, shard)
        return int(shard.size)

    def convert_block_to_ec(self, block_id: str, data_shards: int, parity_shards: int) -> dict:
        """The byte work of the chunkserver-side "EC encode and distribute" command the reference leaves
        as a TODO (metaserver/src/master.rs:2108-2113): verify the stored replicated block, encode it,
        and store every shard as `<block_id>_shard<i>` with its .meta — shards, their .meta arrays, the
        per-shard CRCs (WriteBlock's expected_checksum) and CRC(original) all come from ONE H2D of the block.
        -> {"shard_ids", "shard_checksums", "full_checksum", "original_size"}."""
        from . import erasure
        with open(os.path.join(self.storage_dir, block_id), "rb") as f:
            data = np.frombuffer(f.read(), dtype=np.uint8)
        self.verify_block(block_id, data)
        shards, crcs, full, metas = erasure.encode_with_checksums(data, data_shards, parity_shards, with_meta=True)
        ids = []
        for i, (sh, mt) in enumerate(zip(shards, metas)):
            sid = f"{block_id}_shard{i}"
            with open(os.path.join(self.storage_dir, sid), "wb") as f:
                f.write(sh.tobytes())
            with open(os.path.join(self.storage_dir, sid + ".meta"), "wb") as f:
                f.write(mt.tobytes())
            ids.append(sid)
        return {"shard_ids": ids, "shard_checksums": [int(c) for c in crcs], "full_checksum": int(full),
                "original_size": int(data.size)}

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
