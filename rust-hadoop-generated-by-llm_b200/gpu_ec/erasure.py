"""Host-side mirror of dfs/common/src/erasure.rs (same names, argument meaning, error text).

`encode(data, k, m)` returns k+m shards of ceil(len/k) bytes (erasure.rs:7-24); `decode` fills
every missing slot of the caller's list in place and returns the original bytes truncated to
`original_len` (erasure.rs:30-49); `reconstruct` is the crate call the chunkserver makes directly
(chunkserver.rs:620-623). All compute runs in libgpu_ec.so's CUDA kernels.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np

from ._lib import (ERR_EMPTY_DATA, ERR_ZERO_SHARDS, GpuEcError, check, lib)


class ErasureError(Exception):
    """anyhow::Error equivalent: str(e) carries the reference's message."""


def _wrap(fn):
    def inner(*a, **kw):
        try:
            return fn(*a, **kw)
        except GpuEcError as e:
            raise ErasureError(str(e)) from e
    inner.__name__ = fn.__name__
    inner.__doc__ = fn.__doc__
    return inner


def shard_len(data_len: int, data_shards: int) -> int:
    """erasure.rs:56-59. Panics (AssertionError here) when data_shards == 0."""
    assert data_shards > 0, "data_shards must be > 0"
    return lib().gpu_ec_shard_len(data_len, data_shards)


def _as_u8(data) -> np.ndarray:
    if isinstance(data, np.ndarray):
        return np.ascontiguousarray(data, dtype=np.uint8).reshape(-1)
    return np.frombuffer(bytes(data), dtype=np.uint8)


def _ptr_array(arrs: Sequence[Optional[np.ndarray]]):
    P = (C.c_void_p * len(arrs))()
    for i, a in enumerate(arrs):
        P[i] = a.ctypes.data if a is not None else None
    return P


@_wrap
def encode(data, data_shards: int, parity_shards: int) -> List[np.ndarray]:
    """erasure.rs:7-24 -> list of k+m uint8 arrays, the last `parity_shards` are parity."""
    d = _as_u8(data)
    if data_shards <= 0 or parity_shards <= 0:
        raise GpuEcError(ERR_ZERO_SHARDS)
    if d.size == 0:
        raise GpuEcError(ERR_EMPTY_DATA)
    S = lib().gpu_ec_shard_len(d.size, data_shards)
    n = data_shards + parity_shards
    if n > 256:
        check(lib().gpu_ec_encode(d.ctypes.data, d.size, data_shards, parity_shards, None, None, None, None))
    shards = [np.empty(S, dtype=np.uint8) for _ in range(n)]
    check(lib().gpu_ec_encode(d.ctypes.data, d.size, data_shards, parity_shards, _ptr_array(shards), None, None, None))
    return shards


@_wrap
def encode_with_checksums(data, data_shards: int, parity_shards: int, with_meta: bool = False):
    """The client write path in one call (client/src/mod.rs:326-341,600-611): shards, per-shard
    CRC-32s, CRC-32 of the original buffer, and optionally each shard's big-endian .meta bytes."""
    d = _as_u8(data)
    if data_shards <= 0 or parity_shards <= 0:
        raise GpuEcError(ERR_ZERO_SHARDS)
    if d.size == 0:
        raise GpuEcError(ERR_EMPTY_DATA)
    S = lib().gpu_ec_shard_len(d.size, data_shards)
    n = data_shards + parity_shards
    shards = [np.empty(S, dtype=np.uint8) for _ in range(n)]
    crcs = np.zeros(n, dtype=np.uint32)
    full = C.c_uint32(0)
    metas = [np.empty(4 * ((S + 511) // 512), dtype=np.uint8) for _ in range(n)] if with_meta else None
    check(lib().gpu_ec_encode(d.ctypes.data, d.size, data_shards, parity_shards, _ptr_array(shards),
                              crcs.ctypes.data, C.byref(full), _ptr_array(metas) if metas else None))
    return shards, crcs, full.value, metas


def _prepare(shards: List[Optional[np.ndarray]], n: int):
    if len(shards) != n:
        raise ErasureError("RS reconstruct error: TooFewShards" if len(shards) < n else "RS reconstruct error: TooManyShards")
    present = np.array([0 if s is None else 1 for s in shards], dtype=np.uint8)
    sizes = {int(np.asarray(s).size) for s in shards if s is not None}
    if len(sizes) > 1:
        raise ErasureError("RS reconstruct error: IncorrectShardSize")
    S = sizes.pop() if sizes else 0
    bufs = [np.ascontiguousarray(s, dtype=np.uint8).reshape(-1) if s is not None else np.empty(S, dtype=np.uint8)
            for s in shards]
    return present, S, bufs


@_wrap
def reconstruct(shards: List[Optional[np.ndarray]], data_shards: int, parity_shards: int) -> None:
    """ReedSolomon::new(k,m)?.reconstruct(&mut shards) (chunkserver.rs:620-623): every None slot,
    data or parity, is replaced by its rebuilt shard."""
    n = data_shards + parity_shards
    if data_shards <= 0 or parity_shards <= 0:
        raise GpuEcError(ERR_ZERO_SHARDS)
    present, S, bufs = _prepare(shards, n)
    check(lib().gpu_ec_reconstruct(_ptr_array(bufs), present.ctypes.data, S, data_shards, parity_shards))
    for i in range(n):
        if shards[i] is None:
            shards[i] = bufs[i]


@_wrap
def reconstruct_only(shards: List[Optional[np.ndarray]], data_shards: int, parity_shards: int, want: Sequence[int]) -> None:
    """Rebuild only the listed missing slots (the repair path keeps one shard, chunkserver.rs:626-631);
    other None slots stay None."""
    n = data_shards + parity_shards
    if data_shards <= 0 or parity_shards <= 0:
        raise GpuEcError(ERR_ZERO_SHARDS)
    present, S, bufs = _prepare(shards, n)
    w = (C.c_int * len(want))(*want)
    check(lib().gpu_ec_reconstruct_only(_ptr_array(bufs), present.ctypes.data, S, data_shards, parity_shards, w, len(want)))
    for i in want:
        shards[i] = bufs[i]


@_wrap
def decode(shards: List[Optional[np.ndarray]], data_shards: int, parity_shards: int, original_len: int) -> np.ndarray:
    """erasure.rs:30-49."""
    n = data_shards + parity_shards
    if data_shards <= 0 or parity_shards <= 0:
        raise GpuEcError(ERR_ZERO_SHARDS)
    present, S, bufs = _prepare(shards, n)
    out = np.empty(min(original_len, S * data_shards), dtype=np.uint8)
    check(lib().gpu_ec_decode(_ptr_array(bufs), present.ctypes.data, S, data_shards, parity_shards,
                              original_len, out.ctypes.data if out.size else None))
    for i in range(n):
        if shards[i] is None:
            shards[i] = bufs[i]
    return out


def coding_matrix(k: int, m: int) -> np.ndarray:
    out = np.zeros((k + m, k), dtype=np.uint8)
    check(lib().gpu_ec_rs_matrix(k, m, out.ctypes.data))
    return out


def decode_matrix(k: int, m: int, present):
    pres = np.ascontiguousarray(present, dtype=np.uint8)
    rows = np.zeros((k + m, k), dtype=np.uint8)
    missing = np.zeros(k + m, dtype=np.int32)
    surv = np.zeros(k, dtype=np.int32)
    nm = C.c_int(0)
    check(lib().gpu_ec_rs_decode_matrix(k, m, pres.ctypes.data, rows.ctypes.data, missing.ctypes.data,
                                        surv.ctypes.data, C.byref(nm)))
    return rows[:nm.value].copy(), missing[:nm.value].tolist(), surv.tolist()
