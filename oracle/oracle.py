"""ctypes front end of the CPU ORACLE (test infrastructure, NOT product code).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
import this module. It wraps oracle/libec_oracle.so (ec_oracle.c), the plain-C restatement of
the reference path (dfs/common/src/erasure.rs:7-59, dfs/chunkserver/src/chunkserver.rs:182-292
and the two un-vendored crates). PARITY UNPINNED — see ec_oracle.h.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libec_oracle.so")

POLY_IEEE = 0
POLY_CASTAGNOLI = 1

OK = 0
ERR_ZERO_SHARDS = -1
ERR_EMPTY_DATA = -2
ERR_TOO_MANY_SHARDS = -3
ERR_TOO_FEW_PRESENT = -4
ERR_EMPTY_SHARD = -5
ERR_SINGULAR = -6
ERR_COUNT_MISMATCH = -7
ERR_CRC_MISMATCH = -8


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "ec_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        u8p, u32p, sz = C.c_void_p, C.c_void_p, C.c_size_t
        L.orc_gf_mul.restype = C.c_uint8
        L.orc_gf_mul.argtypes = [C.c_uint8, C.c_uint8]
        L.orc_gf_exp.restype = C.c_uint8
        L.orc_gf_exp.argtypes = [C.c_uint8, C.c_uint]
        L.orc_rs_matrix.argtypes = [C.c_int, C.c_int, u8p]
        L.orc_rs_decode_matrix.argtypes = [C.c_int, C.c_int, u8p, u8p, C.c_void_p, C.c_void_p, C.c_void_p]
        L.orc_shard_len.restype = sz
        L.orc_shard_len.argtypes = [sz, sz]
        L.orc_encode.argtypes = [u8p, sz, C.c_int, C.c_int, u8p]
        L.orc_reconstruct.argtypes = [u8p, u8p, sz, C.c_int, C.c_int]
        L.orc_decode.argtypes = [u8p, u8p, sz, C.c_int, C.c_int, sz, u8p]
        for f in (L.orc_crc32, L.orc_crc32_bytewise):
            f.restype = C.c_uint32
            f.argtypes = [u8p, sz, C.c_int]
        L.orc_crc32_combine.restype = C.c_uint32
        L.orc_crc32_combine.argtypes = [C.c_uint32, C.c_uint32, sz, C.c_int]
        L.orc_crc32_chunks.restype = sz
        L.orc_crc32_chunks.argtypes = [u8p, sz, sz, C.c_int, u32p]
        L.orc_meta_encode_be.argtypes = [u32p, sz, u8p]
        L.orc_verify_chunks.argtypes = [u8p, sz, sz, C.c_int, u8p, sz, C.c_void_p]
        L.orc_fill_synthetic.argtypes = [u8p, sz, C.c_uint64, C.c_uint64]
        L.orc_bench_encode.restype = C.c_double
        L.orc_bench_encode.argtypes = [u8p, sz, sz, C.c_int, C.c_int, C.c_int, C.c_int]
        L.orc_bench_crc_chunks.restype = C.c_double
        L.orc_bench_crc_chunks.argtypes = [u8p, sz, sz, C.c_int, C.c_int]
        L.orc_bench_reconstruct.restype = C.c_double
        L.orc_bench_reconstruct.argtypes = [u8p, u8p, sz, sz, C.c_int, C.c_int, C.c_int]
        _lib = L
    return _lib


def _u8(a) -> np.ndarray:
    if isinstance(a, (bytes, bytearray, memoryview)):
        a = np.frombuffer(bytes(a), dtype=np.uint8)
    return np.ascontiguousarray(a, dtype=np.uint8)


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class OracleError(Exception):
    def __init__(self, code):
        super().__init__(f"oracle error {code}")
        self.code = code


def gf_mul(a, b):
    return lib().orc_gf_mul(a, b)


def gf_exp(a, n):
    return lib().orc_gf_exp(a, n)


def rs_matrix(k, m) -> np.ndarray:
    out = np.zeros((k + m, k), dtype=np.uint8)
    rc = lib().orc_rs_matrix(k, m, _p(out))
    if rc:
        raise OracleError(rc)
    return out


def rs_decode_matrix(k, m, present):
    pres = _u8(np.asarray(present, dtype=np.uint8))
    rows = np.zeros((k + m, k), dtype=np.uint8)
    missing = np.zeros(k + m, dtype=np.int32)
    surv = np.zeros(k, dtype=np.int32)
    nm = C.c_int(0)
    rc = lib().orc_rs_decode_matrix(k, m, _p(pres), _p(rows), _p(missing), _p(surv), C.byref(nm))
    if rc:
        raise OracleError(rc)
    n = nm.value
    return rows[:n].copy(), missing[:n].tolist(), surv.tolist()


def shard_len(n, k):
    return lib().orc_shard_len(n, k)


def encode(data, k, m) -> np.ndarray:
    """-> (k+m, S) uint8 array (erasure.rs:7-24)."""
    d = _u8(data)
    if k <= 0 or m <= 0:
        raise OracleError(ERR_ZERO_SHARDS)
    if d.size == 0:
        raise OracleError(ERR_EMPTY_DATA)
    if k + m > 256:
        raise OracleError(ERR_TOO_MANY_SHARDS)
    S = shard_len(d.size, k)
    out = np.empty((k + m, S), dtype=np.uint8)
    rc = lib().orc_encode(_p(d), d.size, k, m, _p(out))
    if rc:
        raise OracleError(rc)
    return out


def reconstruct(shards: np.ndarray, present, k, m) -> np.ndarray:
    sh = np.ascontiguousarray(shards, dtype=np.uint8).copy()
    pres = _u8(np.asarray(present, dtype=np.uint8))
    rc = lib().orc_reconstruct(_p(sh), _p(pres), sh.shape[1], k, m)
    if rc:
        raise OracleError(rc)
    return sh


def decode(shards: np.ndarray, present, k, m, original_len) -> np.ndarray:
    sh = np.ascontiguousarray(shards, dtype=np.uint8).copy()
    pres = _u8(np.asarray(present, dtype=np.uint8))
    out = np.empty(min(original_len, sh.shape[1] * k), dtype=np.uint8)
    rc = lib().orc_decode(_p(sh), _p(pres), sh.shape[1], k, m, original_len, _p(out))
    if rc:
        raise OracleError(rc)
    return out


def crc32(data, poly=POLY_IEEE) -> int:
    d = _u8(data)
    return lib().orc_crc32(_p(d), d.size, poly)


def crc32_bytewise(data, poly=POLY_IEEE) -> int:
    d = _u8(data)
    return lib().orc_crc32_bytewise(_p(d), d.size, poly)


def crc32_combine(c1, c2, len2, poly=POLY_IEEE) -> int:
    return lib().orc_crc32_combine(c1, c2, len2, poly)


def crc32_chunks(data, chunk=512, poly=POLY_IEEE) -> np.ndarray:
    d = _u8(data)
    n = (d.size + chunk - 1) // chunk
    out = np.empty(n, dtype=np.uint32)
    got = lib().orc_crc32_chunks(_p(d), d.size, chunk, poly, _p(out))
    assert got == n
    return out


def meta_encode_be(crcs) -> bytes:
    c = np.ascontiguousarray(crcs, dtype=np.uint32)
    out = np.empty(4 * c.size, dtype=np.uint8)
    lib().orc_meta_encode_be(_p(c), c.size, _p(out))
    return out.tobytes()


def verify_chunks(data, meta_be, chunk=512, poly=POLY_IEEE):
    """-> (rc, first_bad). rc: 0 | ERR_COUNT_MISMATCH | ERR_CRC_MISMATCH (chunkserver.rs:265-289)."""
    d = _u8(data)
    mb = _u8(meta_be)
    fb = C.c_int64(-1)
    rc = lib().orc_verify_chunks(_p(d), d.size, chunk, poly, _p(mb), mb.size, C.byref(fb))
    return rc, fb.value


def fill_synthetic(n, seed, byte_offset=0) -> np.ndarray:
    out = np.empty(n, dtype=np.uint8)
    lib().orc_fill_synthetic(_p(out), n, seed, byte_offset)
    return out


def bench_encode(data: np.ndarray, stripe_len, n_stripes, k, m, threads, with_crc=False) -> float:
    d = _u8(data)
    assert d.size >= stripe_len * n_stripes
    return lib().orc_bench_encode(_p(d), stripe_len, n_stripes, k, m, threads, int(with_crc))


def bench_crc_chunks(data: np.ndarray, chunk, poly, threads) -> float:
    d = _u8(data)
    return lib().orc_bench_crc_chunks(_p(d), d.size, chunk, poly, threads)


def bench_reconstruct(shards: np.ndarray, present, S, n_stripes, k, m, threads) -> float:
    sh = _u8(shards)
    pres = _u8(np.asarray(present, dtype=np.uint8))
    return lib().orc_bench_reconstruct(_p(sh), _p(pres), S, n_stripes, k, m, threads)
