#!/usr/bin/env python3
"""Plain pinned-memory copy ceiling of this box, no kernels: the denominator for every e2e figure.

One process per GPU (spawned here, started together through a gloo barrier), each timing with CUDA
events on its own GPU:
  h2d        contiguous pinned -> device cudaMemcpyAsync
  d2h        device -> contiguous pinned
  both       h2d and d2h at once on two streams (the encode write path does both)
  h2d_2d     cudaMemcpy2DAsync, source pitch = 11 184 811 bytes (the odd shard_len of a 64 MiB
             RS(6,3) stripe), destination pitch rounded up to 256: what gpu_ec_encode's H2D does
  h2d_2d_al  the same with a 256-byte-aligned source pitch (is the odd pitch the problem?)
Prints one JSON line: per-GPU GB/s (min / median over ranks) and the aggregate.

usage: python tools/pcie_ceiling.py --gpus N [--mib 768] [--iters 20]
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import time


def cudart():
    import torch  # noqa: F401  (loads libcudart into the process)
    import glob
    cands = glob.glob(os.path.join(os.path.dirname(torch.__file__), "lib", "libcudart*.so*"))
    cands += glob.glob("/usr/local/cuda/lib64/libcudart.so*")
    import site
    for sp in site.getsitepackages():
        cands += glob.glob(os.path.join(sp, "nvidia", "cuda_runtime", "lib", "libcudart.so*"))
    for c in cands:
        try:
            return C.CDLL(c)
        except OSError:
            pass
    raise SystemExit("libcudart not found")


def worker(rank, world, mib, iters, port):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    torch.cuda.set_device(rank)
    if world > 1:
        dist.init_process_group("gloo")
    rt = cudart()
    rt.cudaMemcpy2DAsync.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_void_p]
    n = mib << 20
    host = torch.empty(n, dtype=torch.uint8).pin_memory()
    host.fill_(7)
    host2 = torch.empty(n, dtype=torch.uint8).pin_memory()
    dev = torch.empty(n + (1 << 20), dtype=torch.uint8, device="cuda")
    dev2 = torch.empty(n, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    def timed(fn, bytes_per_iter):
        for _ in range(3):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        e0.record(s1)
        for _ in range(iters):
            fn()
        s1.wait_stream(s2)
        e1.record(s1)
        torch.cuda.synchronize()
        wall = time.perf_counter() - t0
        ms = e0.elapsed_time(e1)
        return bytes_per_iter * iters / (ms * 1e-3) / 1e9, bytes_per_iter * iters / wall / 1e9

    def h2d():
        with torch.cuda.stream(s1):
            dev[:n].copy_(host, non_blocking=True)

    def d2h():
        with torch.cuda.stream(s1):
            host2.copy_(dev2, non_blocking=True)

    def both():
        with torch.cuda.stream(s1):
            dev[:n].copy_(host, non_blocking=True)
        with torch.cuda.stream(s2):
            host2.copy_(dev2, non_blocking=True)

    S = 11_184_811
    rows = n // S
    dp = (S + 255) // 256 * 256
    rows = min(rows, (n + (1 << 20)) // dp)

    def mk2d(spitch):
        r = min(rows, n // spitch)

        def f():
            rc = rt.cudaMemcpy2DAsync(dev.data_ptr(), dp, host.data_ptr(), spitch, S, r, 1, s1.cuda_stream)
            if rc:
                raise SystemExit(f"cudaMemcpy2DAsync rc={rc}")
        return f, S * r

    f_odd, b_odd = mk2d(S)
    f_al, b_al = mk2d(dp)
    res = {}
    for name, fn, nb in (("h2d", h2d, n), ("d2h", d2h, n), ("both", both, 2 * n), ("h2d_2d", f_odd, b_odd), ("h2d_2d_al", f_al, b_al)):
        res[name] = timed(fn, nb)[0]
    if world > 1:
        allr = [None] * world
        dist.all_gather_object(allr, res)
    else:
        allr = [res]
    if rank == 0:
        out = {"tool": "pcie_ceiling", "gpus": world, "MiB_per_copy": mib, "iters": iters, "unit": "GB/s per GPU (device clock)"}
        for k in res:
            v = sorted(r[k] for r in allr)
            out[k] = {"min": round(v[0], 2), "median": round(v[len(v) // 2], 2), "sum": round(sum(v), 2)}
        print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--mib", type=int, default=768)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--rank", type=int, default=-1)
    ap.add_argument("--port", type=int, default=29611)
    a = ap.parse_args()
    if a.rank >= 0:
        worker(a.rank, a.gpus, a.mib, a.iters, a.port)
        return
    procs = [subprocess.Popen([sys.executable, __file__, "--gpus", str(a.gpus), "--mib", str(a.mib), "--iters", str(a.iters),
                               "--rank", str(r), "--port", str(a.port)]) for r in range(a.gpus)]
    rc = 0
    for p in procs:
        rc |= p.wait()
    sys.exit(rc)


if __name__ == "__main__":
    main()
