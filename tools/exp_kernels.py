#!/usr/bin/env python3
"""Kernel experiments (not part of the product): time individual device-resident launches.
usage: python tools/exp_kernels.py crc|rs104|rs63 ..."""
import ctypes as C
import os
import signal
import sys

signal.signal(signal.SIGPIPE, signal.SIG_DFL)  # `| head -1` must not raise BrokenPipeError

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "rust-hadoop-generated-by-llm_b200"))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import gpu_ec
from gpu_ec import device
from gpu_ec._lib import check

L = gpu_ec.lib()
PEAK = 6557.8


def timed(fn, steps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def crc():
    n = 1 << 30
    t = torch.empty(n, dtype=torch.uint8, device="cuda")
    device.fill_synthetic(t, 2)
    for chunk in (512, 65536):
        out = torch.empty(n // chunk, dtype=torch.int32, device="cuda")
        ms = timed(lambda: device.crc32_chunks(t, chunk, out=out))
        print(f"crc chunk={chunk}: {n/ms/1e6:.0f} GB/s frac={n/ms/1e6/PEAK:.3f}")
    # whole-buffer CRC of 14 shards of a 64 MiB RS(10,4) stripe (one fused-path group)
    S, pitch = 6710887, 6711040
    o = torch.empty(14, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    ms = timed(lambda: check(L.gpu_ec_crc32_dev(t.data_ptr(), 14, pitch, S, 0, o.data_ptr(), st)), 20)
    print(f"crc whole 14x6.7MB: {ms*1e3:.1f} us  {14*S/ms/1e6:.0f} GB/s")
    ms = timed(lambda: check(L.gpu_ec_crc32_dev(t.data_ptr(), 140, pitch, S, 0, torch.empty(140, dtype=torch.int32, device='cuda').data_ptr(), st)), 10)
    print(f"crc whole 140x6.7MB: {ms*1e3:.1f} us  {140*S/ms/1e6:.0f} GB/s")


def sfold():
    """whole-buffer CRC: sparse-fold kernels vs the table kernel, same buffers"""
    st = torch.cuda.current_stream().cuda_stream
    for n_buf, S in ((140, 6710887), (1400, 6710887), (14, 6710887), (1, 1 << 30), (64, 11184811), (1, 64 << 20)):
        pitch = (S + 255) // 256 * 256
        t = torch.empty(n_buf * pitch, dtype=torch.uint8, device="cuda")
        device.fill_synthetic(t, 2)
        o = torch.empty(n_buf, dtype=torch.int32, device="cuda")
        res = {}
        for mode in ("1", "0"):
            os.environ["GPU_EC_SFOLD"] = "2" if mode == "1" else "0"
            l0 = L.gpu_ec_kernel_launch_count()
            ms = timed(lambda: check(L.gpu_ec_crc32_dev(t.data_ptr(), n_buf, pitch, S, 0, o.data_ptr(), st)), 10)
            res[mode] = (ms, o.cpu().numpy().copy(), (L.gpu_ec_kernel_launch_count() - l0) // 13)
        same = bool((res["1"][1] == res["0"][1]).all())
        for mode, name in (("1", "sfold"), ("0", "table")):
            ms = res[mode][0]
            print(f"crc whole {n_buf} x {S}: {name}: {ms*1e3:.1f} us {n_buf*S/ms/1e6:.0f} GB/s frac={n_buf*S/ms/1e6/PEAK:.3f} launches={res[mode][2]} same={same}")
        del t


def fused(n=256):
    """config 4 slice: RS(10,4) encode + 14 shard CRCs + CRC(original): single pass vs two-pass, and encode alone"""
    S = gpu_ec.erasure.shard_len(64 << 20, 10)
    b = device.StripeBatch(n, 10, 4, S)
    b.fill_synthetic(4, original_len=64 << 20)
    a = 14 * S * n
    for mode, name in (("1", "fused"), ("0", "two-pass")):
        os.environ["GPU_EC_SFOLD"] = mode
        l0 = L.gpu_ec_kernel_launch_count()
        ms = timed(lambda: b.encode_crc(original_len=64 << 20), 5)
        print(f"rs(10,4)+14crc n={n} {name}: {ms:.3f} ms {a/ms/1e6:.0f} GB/s frac={a/ms/1e6/PEAK:.3f} launches/step={(L.gpu_ec_kernel_launch_count()-l0)//8}")
    ms = timed(b.encode, 5)
    print(f"rs(10,4) encode only: {ms:.3f} ms {a/ms/1e6:.0f} GB/s frac={a/ms/1e6/PEAK:.3f}")


def fused63(n=64):
    """RS(6,3) encode + 9 shard CRCs + CRC(original) on 64 MiB blocks: single pass vs two-pass vs encode only"""
    S = gpu_ec.erasure.shard_len(64 << 20, 6)
    b = device.StripeBatch(n, 6, 3, S)
    b.fill_synthetic(4, original_len=64 << 20)
    a = 9 * S * n
    for mode, name in (("1", "fused"), ("0", "two-pass")):
        os.environ["GPU_EC_SFOLD"] = mode
        l0 = L.gpu_ec_kernel_launch_count()
        ms = timed(lambda: b.encode_crc(original_len=64 << 20), 10)
        print(f"rs(6,3)+9crc n={n} {name}: {ms:.3f} ms {a/ms/1e6:.0f} GB/s frac={a/ms/1e6/PEAK:.3f} launches/step={(L.gpu_ec_kernel_launch_count()-l0)//13}")
    ms = timed(b.encode, 10)
    print(f"rs(6,3) encode only: {ms:.3f} ms {a/ms/1e6:.0f} GB/s frac={a/ms/1e6/PEAK:.3f}")


def bigcall():
    """latency of ONE gpu_ec_encode call on a 256 MiB pinned block, RS(6,3), with shard CRCs (set GPU_EC_NO_CHUNKED_COPY=1 to compare)"""
    import time
    import zlib
    from oracle import oracle
    blk, k, m = 256 << 20, 6, 3
    S = (blk + k - 1) // k
    def pinned(nb):
        p = C.c_void_p()
        check(L.gpu_ec_alloc_pinned(C.byref(p), nb))
        return p, np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(nb,))
    ps, src = pinned(blk)
    pp, par = pinned(m * S)
    src[:] = oracle.fill_synthetic(blk, 5)
    ptrs = (C.c_void_p * (k + m))()
    for r in range(m):
        ptrs[k + r] = par.ctypes.data + r * S
    crcs = np.zeros(k + m, np.uint32)
    full = C.c_uint32(0)
    ts = []
    for _ in range(8):
        t0 = time.perf_counter()
        check(L.gpu_ec_encode(src.ctypes.data, blk, k, m, ptrs, crcs.ctypes.data, C.byref(full), None))
        ts.append(time.perf_counter() - t0)
    want = oracle.encode(src, k, m)
    ok = np.array_equal(par.reshape(m, S), want[k:]) and crcs.tolist() == [zlib.crc32(x.tobytes()) for x in want] and full.value == zlib.crc32(src.tobytes())
    print(f"gpu_ec_encode 256 MiB RS(6,3)+CRC, one caller: {min(ts[2:])*1e3:.2f} ms best, {sorted(ts[2:])[len(ts[2:])//2]*1e3:.2f} ms median, "
          f"{blk/min(ts[2:])/1e9:.1f} data GB/s, chunked={'no' if os.environ.get('GPU_EC_NO_CHUNKED_COPY') else 'yes'}, parity_ok={ok}")


def rs(k, m, stripe_len, n):
    S = gpu_ec.erasure.shard_len(stripe_len, k)
    b = device.StripeBatch(n, k, m, S)
    b.fill_synthetic(4)
    ms = timed(b.encode, 5)
    a = (k + m) * S * n
    print(f"rs({k},{m}) n={n} L2_BLOCK_MB={os.environ.get('GPU_EC_L2_BLOCK_MB')} lut4={os.environ.get('GPU_EC_FORCE_LUT4')}: "
          f"{ms:.3f} ms {a/ms/1e6:.0f} GB/s frac={a/ms/1e6/PEAK:.3f}")
    if k == 10:
        # single pass of 5 inputs -> 4 outputs over everything (how fast is one byte-LUT pass?)
        M = gpu_ec.erasure.coding_matrix(k, m)
        rows = np.ascontiguousarray(M[k:, :5])
        ii = (C.c_int * 5)(0, 1, 2, 3, 4)
        oo = (C.c_int * 4)(10, 11, 12, 13)
        d = b.desc()
        st = torch.cuda.current_stream().cuda_stream
        ms = timed(lambda: check(L.gpu_ec_apply_matrix_dev(C.byref(d), rows.ctypes.data, ii, 5, oo, 4, st)), 5)
        a1 = 9 * S * n
        print(f"   one lut8<5> pass (5 in, 4 out) over all stripes: {ms:.3f} ms {a1/ms/1e6:.0f} GB/s (bytes moved)")


if __name__ == "__main__":
    what = sys.argv[1]
    if what == "crc":
        crc()
    elif what == "sfold":
        sfold()
    elif what == "bigcall":
        bigcall()
    elif what == "fused63":
        fused63(int(sys.argv[2]) if len(sys.argv) > 2 else 64)
    elif what == "fused":
        fused(int(sys.argv[2]) if len(sys.argv) > 2 else 256)
    elif what == "rs104":
        rs(10, 4, 64 << 20, int(sys.argv[2]) if len(sys.argv) > 2 else 48)
    elif what == "rs63":
        rs(6, 3, 64 << 20, 64)


def scale():
    """Does RS(10,4) / RS(6,3) throughput depend on the resident size (TLB / DRAM) or on clocks?"""
    import subprocess, threading, time
    for k, m, n in ((10, 4, 48), (10, 4, 1024), (6, 3, 64), (6, 3, 1000)):
        S = gpu_ec.erasure.shard_len(64 << 20, k)
        b = device.StripeBatch(n, k, m, S)
        b.fill_synthetic(4)
        torch.cuda.synchronize()
        proc = subprocess.Popen(["nvidia-smi", "--query-gpu=clocks.sm,power.draw", "--format=csv,noheader,nounits", "-lms", "20"],
                                stdout=subprocess.PIPE, text=True)
        lines = []
        th = threading.Thread(target=lambda: lines.extend(proc.stdout), daemon=True)
        th.start()
        time.sleep(0.1)
        steps = max(5, 400 // max(1, n // 48))
        ms = timed(b.encode, steps, 3)
        time.sleep(0.05)
        proc.terminate()
        th.join(timeout=2)
        clk = sorted(float(x.split(",")[0]) for x in lines if "," in x)
        pw = sorted(float(x.split(",")[1]) for x in lines if "," in x)
        a = (k + m) * S * n
        print(f"rs({k},{m}) n={n}: {a/ms/1e6:.0f} GB/s frac={a/ms/1e6/PEAK:.3f} steps={steps} "
              f"clk min/med/max={clk[0]:.0f}/{clk[len(clk)//2]:.0f}/{clk[-1]:.0f} power med/max={pw[len(pw)//2]:.0f}/{pw[-1]:.0f} W")
        del b
        torch.cuda.empty_cache()


if __name__ == "__main__" and sys.argv[1] == "scale":
    scale()


def reco():
    """reconstruct variants: which kernels do e.g. RS(6,3) 1..3 erasures, RS(4,2), RS(10,4) decode take?"""
    for k, m, missing in ((6, 3, [1, 4, 7]), (6, 3, [2]), (4, 2, [0, 5]), (10, 4, [0, 3, 11, 13]), (5, 5, [0, 1, 2, 3, 4]), (8, 3, [1, 9])):
        S = gpu_ec.erasure.shard_len(8 << 20, k)
        n = 256
        b = device.StripeBatch(n, k, m, S)
        b.fill_synthetic(4)
        b.encode()
        present = [0 if i in missing else 1 for i in range(k + m)]
        ms = timed(lambda: b.reconstruct(present), 10)
        a = (k + len(missing)) * S * n
        print(f"reconstruct rs({k},{m}) missing={missing}: {a/ms/1e6:.0f} GB/s frac={a/ms/1e6/PEAK:.3f}")
        ms = timed(b.encode, 10)
        a = (k + m) * S * n
        print(f"encode      rs({k},{m}): {a/ms/1e6:.0f} GB/s frac={a/ms/1e6/PEAK:.3f}")
        del b
        torch.cuda.empty_cache()


if __name__ == "__main__" and sys.argv[1] == "reco":
    reco()


def verify():
    n = 1 << 30
    t = torch.empty(n, dtype=torch.uint8, device="cuda")
    device.fill_synthetic(t, 2)
    for chunk in (512, 65536):
        out = torch.empty(n // chunk, dtype=torch.int32, device="cuda")
        device.crc32_chunks(t, chunk, out=out, big_endian=True)
        bad = torch.empty(1, dtype=torch.int64, device="cuda")
        ms = timed(lambda: device.verify_chunks(t, chunk, out.view(torch.uint8), out=bad))
        print(f"verify chunk={chunk}: {n/ms/1e6:.0f} GB/s frac={n/ms/1e6/PEAK:.3f} bad={int(bad.item())}")


if __name__ == "__main__" and sys.argv[1] == "verify":
    verify()
