"""ctypes loader for libgpu_ec.so (the C ABI in include/gpu_ec.h).

Fails loudly when the library is missing: there is no Python or CPU fallback for the data path.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO_PATH = os.environ.get("GPU_EC_LIB") or os.path.join(_PKG, "libgpu_ec.so")  # override: kernel experiments only
CSRC = os.path.join(_PKG, "csrc")

OK = 0
ERR_ZERO_SHARDS = -1
ERR_EMPTY_DATA = -2
ERR_TOO_MANY_SHARDS = -3
ERR_TOO_FEW_PRESENT = -4
ERR_EMPTY_SHARD = -5
ERR_SHARD_SIZE = -6
ERR_COUNT_MISMATCH = -7
ERR_CRC_MISMATCH = -8
ERR_CUDA = -9
ERR_OOM = -10
ERR_INVALID_ARG = -11
ERR_NO_DEVICE = -12
ERR_UNALIGNED = -13
ERR_SINGULAR = -14

POLY_IEEE = 0
POLY_CASTAGNOLI = 1


class Stripes(C.Structure):
    """gpu_ec_stripes (include/gpu_ec.h)."""
    _fields_ = [("base", C.c_void_p), ("shard_stride", C.c_size_t), ("stripe_stride", C.c_size_t),
                ("shard_len", C.c_size_t), ("n_stripes", C.c_size_t)]


class GpuEcError(RuntimeError):
    def __init__(self, code: int, detail: str = ""):
        self.code = code
        msg = lib().gpu_ec_strerror(code).decode()
        if detail:
            msg += f": {detail}"
        super().__init__(msg)


def build(force: bool = False) -> str:
    """Compile libgpu_ec.so for sm_100a with nvcc (cross-compiles without a GPU)."""
    srcs = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(os.path.dirname(_PKG), "include", "gpu_ec.h")]
    newest = max(os.path.getmtime(s) for s in srcs)
    if force or not os.path.exists(SO_PATH) or os.path.getmtime(SO_PATH) < newest:
        subprocess.check_call(["make", "-C", CSRC, "-s"])
    return SO_PATH


_lib = None

# name -> (restype, argtypes); every symbol include/gpu_ec.h declares
_vp, _sz, _i, _u32, _u64 = C.c_void_p, C.c_size_t, C.c_int, C.c_uint32, C.c_uint64
SIGNATURES = {
    "gpu_ec_init": (_i, []),
    "gpu_ec_shutdown": (None, []),
    "gpu_ec_device_count": (_i, []),
    "gpu_ec_set_devices": (_i, [_vp, _i]),
    "gpu_ec_strerror": (C.c_char_p, [_i]),
    "gpu_ec_last_error": (C.c_char_p, []),
    "gpu_ec_alloc_pinned": (_i, [C.POINTER(_vp), _sz]),
    "gpu_ec_free_pinned": (_i, [_vp]),
    "gpu_ec_shard_len": (_sz, [_sz, _sz]),
    "gpu_ec_rs_matrix": (_i, [_i, _i, _vp]),
    "gpu_ec_rs_decode_matrix": (_i, [_i, _i, _vp, _vp, _vp, _vp, _vp]),
    "gpu_ec_meta_encode_be": (None, [_vp, _sz, _vp]),
    "gpu_ec_crc32_combine": (_u32, [_u32, _u32, _sz, _i]),
    "gpu_ec_encode": (_i, [_vp, _sz, _i, _i, _vp, _vp, _vp, _vp]),
    "gpu_ec_reconstruct": (_i, [_vp, _vp, _sz, _i, _i]),
    "gpu_ec_reconstruct_only": (_i, [_vp, _vp, _sz, _i, _i, _vp, _i]),
    "gpu_ec_decode": (_i, [_vp, _vp, _sz, _i, _i, _sz, _vp]),
    "gpu_ec_crc32": (_i, [_vp, _sz, _i, _vp]),
    "gpu_ec_crc32_chunks": (_i, [_vp, _sz, _sz, _i, _vp]),
    "gpu_ec_verify_chunks": (_i, [_vp, _sz, _sz, _i, _vp, _sz, _vp]),
    "gpu_ec_block_checksums": (_i, [_vp, _sz, _sz, _i, _vp, _vp]),
    "gpu_ec_block_checksums_batch": (_i, [_sz, _vp, _vp, _sz, _i, _vp, _vp]),
    "gpu_ec_encode_batch": (_i, [_vp, _sz, _sz, _i, _i, _vp, _vp, _vp]),
    "gpu_ec_reconstruct_batch": (_i, [_vp, _vp, _sz, _sz, _i, _i]),
    "gpu_ec_verify_chunks_batch": (_i, [_sz, _vp, _vp, _sz, _i, _vp, _vp, _vp, _vp]),
    "gpu_ec_encode_dev": (_i, [C.POINTER(Stripes), _i, _i, _vp]),
    "gpu_ec_reconstruct_dev": (_i, [C.POINTER(Stripes), _vp, _i, _i, _vp]),
    "gpu_ec_apply_matrix_dev": (_i, [C.POINTER(Stripes), _vp, _vp, _i, _vp, _i, _vp]),
    "gpu_ec_crc32_chunks_dev": (_i, [_vp, _sz, _sz, _sz, _sz, _i, _vp, _sz, _i, _vp]),
    "gpu_ec_crc32_dev": (_i, [_vp, _sz, _sz, _sz, _i, _vp, _vp]),
    "gpu_ec_crc32_dev_noalloc": (_i, [_vp, _sz, _sz, _sz, _i, _vp, _vp]),
    "gpu_ec_verify_chunks_dev": (_i, [_vp, _sz, _sz, _i, _vp, _vp, _vp]),
    "gpu_ec_encode_crc_dev": (_i, [C.POINTER(Stripes), _i, _i, _sz, _i, _vp, _vp, _vp]),
    "gpu_ec_fill_synthetic_dev": (_i, [_vp, _sz, _u64, _u64, _vp]),
    "gpu_ec_kernel_launch_count": (_u64, []),
}


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(SO_PATH):
            raise ImportError(
                f"{SO_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(gpu_ec has no CPU fallback)")
        L = C.CDLL(SO_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError here = header/library drift: fail loudly
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def check(rc: int) -> None:
    if rc != OK:
        detail = lib().gpu_ec_last_error().decode() if rc in (ERR_CUDA, ERR_OOM, ERR_NO_DEVICE, ERR_INVALID_ARG) else ""
        raise GpuEcError(rc, detail)
