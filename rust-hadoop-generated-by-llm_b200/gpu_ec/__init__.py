"""gpu_ec — Python mirror of the reference's Rust interface for the RS + CRC byte path, on top of
the C ABI of libgpu_ec.so (hand-written CUDA for B200 / sm_100a; no CPU fallback).

    erasure.encode / decode / shard_len     <- dfs/common/src/erasure.rs:7,30,56
    erasure.reconstruct                     <- ReedSolomon::reconstruct, chunkserver.rs:620-623
    checksum.calculate_checksums / verify_block / verify_partial_read / crc32
                                            <- dfs/chunkserver/src/chunkserver.rs:182-351,746-750
    client.prepare_ec_write / read_ec_block <- dfs/client/src/mod.rs:600-611, 1131-1164 (no RPCs)
    device.*                                <- device-resident batches (torch tensors), used by bench.py
"""
from . import _lib
from ._lib import GpuEcError, POLY_CASTAGNOLI, POLY_IEEE, build, lib  # noqa: F401
from . import erasure, checksum, client  # noqa: F401,E402

__all__ = ["erasure", "checksum", "client", "GpuEcError", "POLY_IEEE", "POLY_CASTAGNOLI", "build", "lib"]
