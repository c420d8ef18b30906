"""The byte-path part of the client's EC write / read orchestration (dfs/client/src/mod.rs), without
the RPCs: what `create_file_from_buffer_ec` (:496-677) computes before it sends k+m WriteBlocks, and
what `read_ec_block` (:1110-1165) does with the shards it fetched. Compute runs in libgpu_ec.so."""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import numpy as np

from . import erasure


@dataclass
class EcWritePlan:
    """One WriteBlockRequest per shard (data=shard, expected_checksum_crc32c, shard_index) plus the
    BlockChecksumInfo of CompleteFile (mod.rs:613-656)."""
    shards: List[np.ndarray]
    shard_checksums: List[int]
    full_checksum: int      # BlockChecksumInfo.checksum_crc32c = CRC-32 of the original buffer
    actual_size: int        # BlockChecksumInfo.actual_size = original length
    data_shards: int
    parity_shards: int


def prepare_ec_write(buffer, data_shards: int, parity_shards: int) -> EcWritePlan:
    """mod.rs:600-611: encode, then CRC each of the k+m shards, then CRC the original buffer —
    three passes over the bytes in the reference, one H2D here."""
    shards, crcs, full, _ = erasure.encode_with_checksums(buffer, data_shards, parity_shards)
    n = int(np.asarray(buffer).size if isinstance(buffer, np.ndarray) else len(buffer))
    return EcWritePlan(shards, [int(c) for c in crcs], int(full), n, data_shards, parity_shards)


def read_ec_block(block_id: str, opt_shards: List[Optional[np.ndarray]], data_shards: int, parity_shards: int,
                  original_size: int) -> np.ndarray:
    """mod.rs:1131-1164 after the fetch: bail when fewer than k shards arrived, concatenate when all
    data shards are present (no RS), else degraded read through erasure.decode."""
    total = data_shards + parity_shards
    opt = list(opt_shards[:total]) + [None] * max(0, total - len(opt_shards))
    available = sum(s is not None for s in opt)
    if available < data_shards:
        raise erasure.ErasureError(
            f"EC block {block_id} unrecoverable: only {available}/{data_shards} shards available")
    if all(s is not None for s in opt[:data_shards]):
        out = np.concatenate([np.asarray(s, dtype=np.uint8).reshape(-1) for s in opt[:data_shards]])
        return out[:original_size]
    return erasure.decode(opt, data_shards, parity_shards, original_size)
