#!/usr/bin/env python3
"""bench.py — RS(6,3) encode / reconstruct + CRC-32 throughput on B200 (the BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--no-sub]

One JSON line on stdout (rank 0). A "step" is one pass of the hot path over one device-resident
batch: RS(6,3) encode of 64 blocks of 64 MiB per GPU (BASELINE config 1 batched as SURVEY.md §8d
prescribes; 6.4 GB per GPU, far larger than the 126 MB L2, so no L2 flush is needed between steps).
`value` counts ALGORITHMIC bytes, (k+m)/k x data, as north_star asks. `e2e` is the same metric
through the host-pointer C ABI (gpu_ec_encode_batch) with pinned host buffers, H2D/D2H inside the
timed region. `sub` carries the other configs (reconstruct, CRC 512 B / 64 KiB, fused RS(10,4)+CRC),
each with its own roofline fraction. Stripes are independent: rank r owns its own stripes, no
collective touches the data path (weak scaling).

--impl reference times the CPU oracle (C port of the reference algorithm; the Rust crates cannot
be built here) on all host cores, on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "rust-hadoop-generated-by-llm_b200"))

K_DATA, M_PAR = 6, 3
STRIPE_LEN = 64 << 20
STRIPES_PER_GPU = 64
SEED = 0xEC000001
METRIC = "rs63_encode_algorithmic_GBps"
UNIT = "GB/s"
# Plain pinned cudaMemcpyAsync on this pool's boxes, no kernels, N processes at once (tools/pcie_ceiling.py;
# profiles/r02/r02_pcie_{1,2,4,8}.json): aggregate GB/s over the N GPUs for H2D alone and for H2D + D2H together.
# The VM's host side, not the library, sets the e2e scaling curve: 8 GPUs together get 225 GB/s H2D but only
# 151 GB/s when both directions run, 4 GPUs 115 / 102.
PCIE_CEILING = {1: {"h2d": 55.6, "both": 91.5}, 2: {"h2d": 111.2, "both": 153.5}, 4: {"h2d": 115.3, "both": 101.8}, 8: {"h2d": 225.5, "both": 150.7}}


def peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=lambda: self.lines.extend(self.proc.stdout), daemon=True)
            self.t.start()
        except Exception:  # noqa: BLE001
            self.proc = None

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, reasons = [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def make_config(n_local, S):
    """The workload description both arms print (identical keys and values: same_config)."""
    return {"workload": "rs63_encode_64MiB_blocks", "k": K_DATA, "m": M_PAR, "stripe_len": STRIPE_LEN, "shard_len": S,
            "stripes_per_gpu": n_local, "partition": "stripe i -> gpu i mod G, no collective",
            "l2": "inputs (6.4 GB/GPU) exceed the 126 MB L2; no flush needed", "bytes_counted": "(k+m)*shard_len per stripe"}


def cpu_reference(steps, warmup, full_line, n_gpus=1, n_local=STRIPES_PER_GPU):
    """The reference arm / cpu_baseline: C oracle (port) on all host cores, bounded sample of the workload:
    16 of its 64 MiB stripes per pass, one stripe per thread, `warmup` untimed passes then `steps` timed ones."""
    from oracle import oracle
    threads = os.cpu_count() or 1
    n_stripes = max(4, min(threads, 16))
    data = oracle.fill_synthetic(STRIPE_LEN * n_stripes, SEED)
    S = oracle.shard_len(STRIPE_LEN, K_DATA)
    alg = (K_DATA + M_PAR) * S * n_stripes
    for _ in range(warmup):
        oracle.bench_encode(data, STRIPE_LEN, n_stripes, K_DATA, M_PAR, threads)
    ts = [oracle.bench_encode(data, STRIPE_LEN, n_stripes, K_DATA, M_PAR, threads) for _ in range(max(1, steps))]
    t = sum(ts) / len(ts)
    val = alg / t / 1e9
    t1 = min(oracle.bench_encode(data, STRIPE_LEN, 1, K_DATA, M_PAR, 1) for _ in range(2))  # ONE reference call: it is single-threaded
    base = {"value": round(val, 3), "unit": UNIT, "cores": threads, "kind": "port",
            "single_thread_value": round((K_DATA + M_PAR) * S / t1 / 1e9, 3),
            "sample": f"{n_stripes} of the 64 MiB RS(6,3) stripes per pass, one stripe per thread, {warmup} warm + {len(ts)} timed pass(es); "
                      "C restatement of reed-solomon-erasure 6.0.0's scalar table path incl. both input copies of erasure.rs:17-19"}
    if not full_line:
        return base
    return {"metric": METRIC, "value": round(val, 3), "unit": UNIT, "impl": "reference", "n_gpus": n_gpus, "steps": len(ts),
            "warmup": warmup, "ms_per_step": round(t * 1e3, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic (splitmix64 counter stream)",
            "config": make_config(n_local, S),
            "cpu_baseline": base,
            "e2e": {"value": round(val, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


def cpu_others(oracle, np):
    """CPU side of the other BASELINE configs on bounded samples: C port at all cores and at one thread, plus zlib's
    crc32() as the independent CRC row. -> {workload: {...}}"""
    import zlib
    from concurrent.futures import ThreadPoolExecutor
    threads = os.cpu_count() or 1
    out = {}
    # config 3: RS(6,3) reconstruct of shards {1,4,7} over 1 MiB stripes (chunkserver.rs:620-623)
    k, m, L1 = 6, 3, 1 << 20
    S3 = oracle.shard_len(L1, k)
    n3 = 256
    sh = np.empty((n3, k + m, S3), np.uint8)
    d = oracle.fill_synthetic(L1 * 8, 0xEC000003)
    for i in range(n3):
        sh[i] = oracle.encode(d[(i % 8) * L1:(i % 8 + 1) * L1], k, m)
    pres = [1, 0, 1, 1, 0, 1, 1, 0, 1]
    oracle.bench_reconstruct(sh, pres, S3, n3, k, m, threads)
    tN = min(oracle.bench_reconstruct(sh, pres, S3, n3, k, m, threads) for _ in range(3))
    t1 = oracle.bench_reconstruct(sh[:16], pres, S3, 16, k, m, 1)
    out["rs63_reconstruct_3_erasures_4096x1MiB"] = {"cpu_port_GBps": round(9 * S3 * n3 / tN / 1e9, 2), "cpu_port_1thread_GBps": round(9 * S3 * 16 / t1 / 1e9, 3),
                                                    "cpu_cores": threads, "cpu_sample": f"{n3} of the 4096 stripes, 3 passes (best)"}
    del sh
    # config 4: RS(10,4) encode + 14 shard CRCs + CRC(original) (client/src/mod.rs:600-611), 64 MiB stripes
    k, m = 10, 4
    S4 = oracle.shard_len(STRIPE_LEN, k)
    n4 = max(4, min(threads, 16))
    d = oracle.fill_synthetic(STRIPE_LEN * n4, 0xEC000004)
    oracle.bench_encode(d, STRIPE_LEN, n4, k, m, threads, True)
    tN = min(oracle.bench_encode(d, STRIPE_LEN, n4, k, m, threads, True) for _ in range(3))
    t1 = oracle.bench_encode(d, STRIPE_LEN, 1, k, m, 1, True)
    tNe = min(oracle.bench_encode(d, STRIPE_LEN, n4, k, m, threads, False) for _ in range(2))
    out["rs104_encode_plus_14_crc_64MiB_stripes"] = {"cpu_port_GBps": round(14 * S4 * n4 / tN / 1e9, 2), "cpu_port_1thread_GBps": round(14 * S4 / t1 / 1e9, 3),
                                                     "cpu_cores": threads, "cpu_sample": f"{n4} of the 1024 stripes, 3 passes (best)"}
    out["rs104_encode_only"] = {"cpu_port_GBps": round(14 * S4 * n4 / tNe / 1e9, 2), "cpu_cores": threads}
    S6 = oracle.shard_len(STRIPE_LEN, K_DATA)
    tN = min(oracle.bench_encode(d, STRIPE_LEN, n4, K_DATA, M_PAR, threads, True) for _ in range(3))
    out["rs63_encode_plus_9_crc_64MiB_blocks"] = {"cpu_port_GBps": round(9 * S6 * n4 / tN / 1e9, 2), "cpu_cores": threads,
                                                  "cpu_sample": f"{n4} of the 64 blocks, 3 passes (best)"}
    # config 2: CRC-32 IEEE per 512 B and per 64 KiB: PCLMULQDQ port (what crc32fast runs on this CPU) and zlib
    samp = d[:256 << 20]
    for chunk in (512, 65536):
        oracle.bench_crc_chunks(samp, chunk, 0, threads)
        tN = min(oracle.bench_crc_chunks(samp, chunk, 0, threads) for _ in range(3))
        t1 = oracle.bench_crc_chunks(samp[:64 << 20], chunk, 0, 1)
        out[f"crc32_ieee_{chunk}B_chunks_1GiB"] = {"cpu_port_GBps": round(samp.size / tN / 1e9, 2), "cpu_port_1thread_GBps": round((64 << 20) / t1 / 1e9, 2),
                                                  "cpu_cores": threads, "cpu_sample": "256 MiB of the 1 GiB, 3 passes (best)"}
    mv = memoryview(samp)
    step = 64 << 10
    t0 = time.perf_counter()
    for o in range(0, 64 << 20, step):
        zlib.crc32(mv[o:o + step])
    tz1 = time.perf_counter() - t0
    slabs = [mv[i * (samp.size // threads):(i + 1) * (samp.size // threads)] for i in range(threads)]
    def zl(sl):
        for o in range(0, len(sl), step):
            zlib.crc32(sl[o:o + step])
    with ThreadPoolExecutor(threads) as ex:
        t0 = time.perf_counter()
        list(ex.map(zl, slabs))
        tzN = time.perf_counter() - t0
    out["crc32_ieee_65536B_chunks_1GiB"].update({"cpu_zlib_GBps": round(samp.size / tzN / 1e9, 2), "cpu_zlib_1thread_GBps": round((64 << 20) / tz1 / 1e9, 2),
                                                 "cpu_zlib_note": "zlib.crc32 per 64 KiB chunk (releases the GIL): the independent CRC row"})
    return out


def ncu_traffic(kernel_substr, alg_bytes):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel_substr` whose algorithmic bytes equal
    alg_bytes, from the committed ncu --set full summaries (profiles/ncu_traffic.json); None when no capture of that
    kernel and shape is committed (the value is then not measured by anything in the repo)."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            for e in json.load(f)["captures"]:
                if kernel_substr in e["kernel"] and int(e["algorithmic_bytes_per_launch"]) == int(alg_bytes):
                    return int(e["dram_bytes_read"]) + int(e["dram_bytes_write"]), e.get("source")
    except Exception:  # noqa: BLE001
        pass
    return None, None


def bind_to_gpu_numa_node(torch, dev):
    """Run this rank on the CPUs local to its GPU's PCIe root (sysfs local_cpulist), so the pinned
    staging buffers are first-touched on the NUMA node the DMA engine reads from. Without it the
    e2e figure at 8 ranks is bounded by the inter-socket link, not by PCIe. Returns a description."""
    try:
        pr = torch.cuda.get_device_properties(dev)
        bdf = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        with open(f"/sys/bus/pci/devices/{bdf}/local_cpulist") as f:
            spec = f.read().strip()
        cpus = set()
        for part in spec.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as f:
                node = f.read().strip()
            return f"gpu {bdf}: numa node {node}, {len(cpus)} local cpus"
    except Exception as e:  # noqa: BLE001
        return f"not bound ({type(e).__name__})"
    return "not bound"


def run_e2e(L, oracle, np, C, check, barrier, partition, dist, world, rank, S, n_e2e):
    hin, hout = C.c_void_p(), C.c_void_p()
    check(L.gpu_ec_alloc_pinned(C.byref(hin), n_e2e * STRIPE_LEN))
    check(L.gpu_ec_alloc_pinned(C.byref(hout), n_e2e * M_PAR * S))
    src = np.ctypeslib.as_array(C.cast(hin, C.POINTER(C.c_uint8)), shape=(n_e2e * STRIPE_LEN,))
    src[:] = oracle.fill_synthetic(n_e2e * STRIPE_LEN, SEED + 100 + rank)

    def e2e_step():
        check(L.gpu_ec_encode_batch(hin, STRIPE_LEN, n_e2e, K_DATA, M_PAR, hout, None, None))

    for _ in range(3):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    e2e_steps = 20
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_ms = (time.perf_counter() - t0) * 1e3 / e2e_steps
    e2e_ms, _ = partition.aggregate(e2e_ms, 0.0, dist if world > 1 else None)
    e2e_val = (K_DATA + M_PAR) * S * n_e2e * world / (e2e_ms * 1e-3) / 1e9
    par = np.ctypeslib.as_array(C.cast(hout, C.POINTER(C.c_uint8)), shape=(n_e2e, M_PAR, S))
    want = oracle.encode(src[:STRIPE_LEN], K_DATA, M_PAR)
    if not np.array_equal(par[0], want[K_DATA:]):
        raise SystemExit("PARITY FAILURE on the e2e path")
    L.gpu_ec_free_pinned(hin)
    L.gpu_ec_free_pinned(hout)

    return e2e_val, e2e_ms, e2e_steps


def run_config5(L, oracle, np, C, check, rank):
    """BASELINE config 5: 32 concurrent client streams over 256 MiB blocks, end to end through the
    host-pointer C ABI with pinned buffers. write = H2D data -> RS(6,3) encode + 9 shard CRCs +
    CRC(original) -> D2H parity + CRCs (gpu_ec_encode); read = H2D block + .meta -> per-512-byte
    verify -> D2H verdict (gpu_ec_verify_chunks). Wall clock over all threads, data bytes counted."""
    import threading
    import zlib
    blk, k, m, n_thr, n_src, iters = 256 << 20, K_DATA, M_PAR, 32, 4, 2
    S = (blk + k - 1) // k
    def pinned(nbytes):
        p = C.c_void_p()
        check(L.gpu_ec_alloc_pinned(C.byref(p), nbytes))
        return p, np.ctypeslib.as_array(C.cast(p, C.POINTER(C.c_uint8)), shape=(nbytes,))
    srcs = [pinned(blk) for _ in range(n_src)]
    for i, (_, a) in enumerate(srcs):
        a[:] = oracle.fill_synthetic(blk, 0xEC000005 + 16 * rank + i)
    pars = [pinned(m * S) for _ in range(n_thr)]
    crcs = [np.zeros(k + m, np.uint32) for _ in range(n_thr)]
    fulls = [C.c_uint32(0) for _ in range(n_thr)]
    metas = []
    for _, a in srcs:  # .meta sidecars for the read side, produced by the library itself
        out = np.zeros(blk // 512, np.uint32)
        check(L.gpu_ec_crc32_chunks(a.ctypes.data, blk, 512, 0, out.ctypes.data))
        mb = np.zeros(4 * out.size, np.uint8)
        L.gpu_ec_meta_encode_be(out.ctypes.data, out.size, mb.ctypes.data)
        metas.append(mb)
    errs = []
    def writer(t):
        try:
            src = srcs[t % n_src][1]
            ptrs = (C.c_void_p * (k + m))()
            for r in range(m):
                ptrs[k + r] = pars[t][1].ctypes.data + r * S
            for _ in range(iters):
                check(L.gpu_ec_encode(src.ctypes.data, blk, k, m, ptrs, crcs[t].ctypes.data, C.byref(fulls[t]), None))
        except Exception as e:  # noqa: BLE001
            errs.append(repr(e))
    def reader(t):
        try:
            src, mb = srcs[t % n_src][1], metas[t % n_src]
            bad = C.c_int64(-1)
            for _ in range(iters):
                check(L.gpu_ec_verify_chunks(src.ctypes.data, blk, 512, 0, mb.ctypes.data, mb.size, C.byref(bad)))
                if bad.value != -1:
                    raise RuntimeError("verify reported corruption")
        except Exception as e:  # noqa: BLE001
            errs.append(repr(e))
    # degraded read (client/src/mod.rs:1143-1164): shards {1,4,7} (2 data + 1 parity) did not arrive; the k first
    # survivors go H2D, one pass rebuilds the three, they come back D2H (gpu_ec_reconstruct)
    want0 = oracle.encode(srcs[0][1], k, m)
    n_deg = 4  # shard sets (pinned, 9 x 44.7 MB each); threads share them read-only except the rebuilt slots
    deg = [pinned((k + m) * S) for _ in range(n_deg)]
    for _, a in deg:
        a[:] = want0.reshape(-1)
    present = np.array([1, 0, 1, 1, 0, 1, 1, 0, 1], np.uint8)
    def degraded(t):
        try:
            base = deg[t % n_deg][1]
            # every thread rebuilds into its own scratch rows (the parity buffers of the write leg are free now)
            ptrs = (C.c_void_p * (k + m))()
            for i in range(k + m):
                ptrs[i] = base.ctypes.data + i * S
            for r, idx in enumerate((1, 4, 7)):
                ptrs[idx] = pars[t][1].ctypes.data + r * S
            for _ in range(iters):
                check(L.gpu_ec_reconstruct(ptrs, present.ctypes.data, S, k, m))
        except Exception as e:  # noqa: BLE001
            errs.append(repr(e))
    def run(fn):
        th = [threading.Thread(target=fn, args=(t,)) for t in range(n_thr)]
        t0 = time.perf_counter()
        [x.start() for x in th]
        [x.join() for x in th]
        return time.perf_counter() - t0
    run(writer)  # warm-up: staging buffers grow to size
    tw = run(writer)
    run(reader)
    tr = run(reader)
    if errs:
        raise SystemExit("config 5 failed: " + errs[0])
    want = want0  # parity gate on one block (thread 0 used source 0)
    got = pars[0][1].reshape(m, S)
    ok = np.array_equal(got, want[k:]) and crcs[0].tolist() == [zlib.crc32(x.tobytes()) for x in want]
    ok = ok and fulls[0].value == zlib.crc32(srcs[0][1].tobytes())
    if not ok:
        raise SystemExit("PARITY FAILURE config 5")
    run(degraded)
    td = run(degraded)
    if errs:
        raise SystemExit("config 5 degraded read failed: " + errs[0])
    rebuilt = pars[5][1].reshape(m, S)
    if not (np.array_equal(rebuilt[0], want0[1]) and np.array_equal(rebuilt[1], want0[4]) and np.array_equal(rebuilt[2], want0[7])):
        raise SystemExit("PARITY FAILURE config 5 degraded read")
    for p, _ in srcs + pars + deg:
        L.gpu_ec_free_pinned(p)
    total = n_thr * iters * blk
    return {"threads": n_thr, "block_bytes": blk, "blocks_per_thread": iters,
            "write_data_GBps": round(total / tw / 1e9, 2), "read_verify_data_GBps": round(total / tr / 1e9, 2),
            "degraded_read_data_GBps": round(n_thr * iters * k * S / td / 1e9, 2),
            "note": "per GPU, wall clock, PCIe-bound: write moves 1.0x data H2D + 0.5x D2H, read-verify 1.0x H2D, "
                    "degraded read (shards 1,4,7 rebuilt) 1.0x H2D + 0.5x D2H; data bytes = k*shard_len per block"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=500)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-sub", action="store_true", help="skip the secondary configs")
    ap.add_argument("--stripes", type=int, default=STRIPES_PER_GPU)
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-pointer leg (profiling runs)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))

    if args.impl == "reference":
        if rank == 0:
            print(json.dumps(cpu_reference(min(args.steps, 20), max(3, args.warmup), True, args.gpus, args.stripes)), flush=True)
        return 0

    import numpy as np
    import torch
    import torch.distributed as dist
    import gpu_ec
    from gpu_ec import device, partition
    from gpu_ec._lib import check
    import ctypes as C

    warmup = max(3, args.warmup)
    torch.cuda.set_device(local)
    numa = bind_to_gpu_numa_node(torch, local)  # before any pinned allocation: first touch decides the node
    if world > 1:
        # bookkeeping only (barrier, max elapsed): stripes are independent, so the data path has no
        # collective and NCCL/NVLink stay unused by design; gloo keeps stdout to the one JSON line
        dist.init_process_group("gloo")
    L = gpu_ec.lib()
    devs = (C.c_int * 1)(local)
    check(L.gpu_ec_set_devices(devs, 1))
    check(L.gpu_ec_init())

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps, warm):
        """-> (ms per step as max over ranks, launches in the timed region)."""
        for _ in range(warm):
            fn()
        barrier()
        l0 = L.gpu_ec_kernel_launch_count()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1) / steps
        ms, _ = partition.aggregate(ms, 0.0, dist if world > 1 else None)
        return ms, (L.gpu_ec_kernel_launch_count() - l0)

    peak, peak_src = peaks()
    S = gpu_ec.erasure.shard_len(STRIPE_LEN, K_DATA)
    n_local = args.stripes  # weak scaling: every rank owns `stripes` stripes of the global round-robin
    batch = device.StripeBatch(n_local, K_DATA, M_PAR, S)
    # global stripe id of local stripe i is i*world + rank; seeds differ per rank so ranks hold different bytes
    batch.fill_synthetic(SEED + rank, original_len=STRIPE_LEN)
    torch.cuda.synchronize()

    # ---- parity gate: no number is reported unless the bytes match the oracle
    from oracle import oracle
    batch.encode()
    torch.cuda.synchronize()
    for i in (0, n_local - 1):
        data = np.concatenate([batch.shard_host(i, j) for j in range(K_DATA)])
        want = oracle.encode(data, K_DATA, M_PAR)
        for s in range(K_DATA, K_DATA + M_PAR):
            if not np.array_equal(batch.shard_host(i, s), want[s]):
                raise SystemExit(f"PARITY FAILURE stripe {i} shard {s}: refusing to report a number")
    batch.buf[:, K_DATA:] = 0

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms, launches = timed(batch.encode, args.steps, warmup)
    clocks = sampler.stop() if rank == 0 else None
    alg_bytes_rank = (K_DATA + M_PAR) * S * n_local
    value = alg_bytes_rank * world / (ms * 1e-3) / 1e9
    per_gpu = alg_bytes_rank / (ms * 1e-3) / 1e9

    # ---- e2e: host-pointer C ABI, pinned buffers, H2D + D2H inside the timed region
    n_e2e = 16
    e2e_val = e2e_ms = e2e_steps = None
    if not args.no_e2e:
        e2e_val, e2e_ms, e2e_steps = run_e2e(L, oracle, np, C, check, barrier, partition, dist, world, rank, S, n_e2e)
    # ---- secondary configs (device-resident), each with its own roofline fraction
    sub = {}
    if not args.no_sub:
        # the headline code with its checksums (client/src/mod.rs:600-611 at RS(6,3)): parity + 9 shard CRCs + CRC(original), one pass
        import zlib
        crc63, full63 = batch.encode_crc(original_len=STRIPE_LEN)
        torch.cuda.synchronize()
        d0 = np.concatenate([batch.shard_host(0, j) for j in range(K_DATA)])
        w0 = oracle.encode(d0[:STRIPE_LEN], K_DATA, M_PAR)
        if device.u32(crc63)[0].tolist() != [zlib.crc32(x.tobytes()) for x in w0] or int(device.u32(full63)[0]) != zlib.crc32(d0[:STRIPE_LEN].tobytes()):
            raise SystemExit("PARITY FAILURE fused RS(6,3)+CRC")
        ms63, l63 = timed(lambda: batch.encode_crc(original_len=STRIPE_LEN), 10, 3)
        sub["rs63_encode_plus_9_crc_64MiB_blocks"] = {"GBps_algorithmic": round(alg_bytes_rank * world / ms63 / 1e6, 1), "ms": round(ms63, 3),
                                                      "roofline_frac": round(alg_bytes_rank / ms63 / 1e6 / peak, 4), "launches_per_step": int(l63) // 10,
                                                      "note": "one pass over HBM: single-CTA byte-LUT encode with the 9 shard CRCs accumulated in registers, then the finisher"}
        del batch
        torch.cuda.empty_cache()
        # config 3: 4096 stripes of 1 MiB, RS(6,3), shards {1,4,7} erased -> one pass, reads k writes e
        S3 = gpu_ec.erasure.shard_len(1 << 20, 6)
        b3 = device.StripeBatch(4096, 6, 3, S3)
        b3.fill_synthetic(0xEC000003 + rank)
        b3.encode()
        torch.cuda.synchronize()
        gate = sorted({0, 1, 2, 3, 2047, 2048, 4094, 4095})  # 8 of the 4096 stripes, oracle-encoded from their data shards
        want3 = {i: oracle.encode(np.concatenate([b3.shard_host(i, j) for j in range(6)]), 6, 3) for i in gate}
        b3.buf[:, [1, 4, 7]] = 0
        present = [1, 0, 1, 1, 0, 1, 1, 0, 1]
        b3.reconstruct(present)
        torch.cuda.synchronize()
        for i in gate:
            if not all(np.array_equal(b3.shard_host(i, s_), want3[i][s_]) for s_ in range(9)):
                raise SystemExit(f"PARITY FAILURE reconstruct (stripe {i} differs from the oracle)")
        ms3, _ = timed(lambda: b3.reconstruct(present), 10, 3)
        a3 = 9 * S3 * 4096
        sub["rs63_reconstruct_3_erasures_4096x1MiB"] = {"GBps_algorithmic": round(a3 * world / ms3 / 1e6, 1), "ms": round(ms3, 4),
                                                       "roofline_frac": round(a3 / ms3 / 1e6 / peak, 4)}
        del b3
        torch.cuda.empty_cache()
        # config 2: 1 GiB, CRC-32 IEEE per 512 B (the .meta layout) and per 64 KiB, write side + verify side
        n2 = 1 << 30
        t2 = torch.empty(n2, dtype=torch.uint8, device="cuda")
        device.fill_synthetic(t2, 0xEC000002 + rank)
        for chunk in (512, 65536):
            out = torch.empty(n2 // chunk, dtype=torch.int32, device="cuda")
            msw, _ = timed(lambda: device.crc32_chunks(t2, chunk, out=out, big_endian=True), 10, 3)
            bad = torch.empty(1, dtype=torch.int64, device="cuda")
            meta = out.view(torch.uint8)
            msv, _ = timed(lambda: device.verify_chunks(t2, chunk, meta, out=bad), 10, 3)
            if int(bad.item()) != -1:
                raise SystemExit("PARITY FAILURE crc verify")
            a2 = n2 + 4 * (n2 // chunk)
            sub[f"crc32_ieee_{chunk}B_chunks_1GiB"] = {
                "write_GBps": round(a2 * world / msw / 1e6, 1), "verify_GBps": round(a2 * world / msv / 1e6, 1),
                "roofline_frac_write": round(a2 / msw / 1e6 / peak, 4), "roofline_frac_verify": round(a2 / msv / 1e6 / peak, 4)}
        host_sample = t2[:64 << 20].cpu().numpy()   # 64 MiB slice of the timed buffer, both chunk sizes, against the oracle
        for chunk in (512, 65536):
            chk = device.u32(device.crc32_chunks(t2, chunk))[:(64 << 20) // chunk]
            if not np.array_equal(chk, oracle.crc32_chunks(host_sample, chunk)):
                raise SystemExit(f"PARITY FAILURE crc chunks ({chunk} B) vs oracle")
        del host_sample
        outc = torch.empty(n2 // 512, dtype=torch.int32, device="cuda")
        msc, _ = timed(lambda: device.crc32_chunks(t2, 512, poly=gpu_ec.POLY_CASTAGNOLI, out=outc), 5, 3)
        sub["crc32c_castagnoli_512B_chunks_1GiB"] = {"write_GBps": round((n2 + 4 * (n2 // 512)) * world / msc / 1e6, 1),
                                                    "note": "extra variant; the reference computes IEEE"}
        del t2, out, outc
        torch.cuda.empty_cache()
        # config 4 (per-GPU slice): fused RS(10,4) encode + 14 shard CRCs + CRC(original), 64 MiB stripes
        S4 = gpu_ec.erasure.shard_len(64 << 20, 10)
        n4 = max(1, 1024 // world)  # config 4: 64 GiB of 64 MiB stripes in total (96 GB resident at 1 GPU), split over the ranks
        b4 = device.StripeBatch(n4, 10, 4, S4)
        b4.fill_synthetic(0xEC000004 + rank, original_len=64 << 20)
        crc4, full4 = b4.encode_crc(original_len=64 << 20)
        torch.cuda.synchronize()
        d0 = np.concatenate([b4.shard_host(0, j) for j in range(10)])
        w0 = oracle.encode(d0, 10, 4)
        import zlib
        ok = all(np.array_equal(b4.shard_host(0, s), w0[s]) for s in range(10, 14))
        ok = ok and device.u32(crc4)[0].tolist() == [zlib.crc32(x.tobytes()) for x in w0]
        ok = ok and int(device.u32(full4)[0]) == zlib.crc32(d0[:64 << 20].tobytes())
        if not ok:
            raise SystemExit("PARITY FAILURE fused RS(10,4)+CRC")
        ms4, l4 = timed(lambda: b4.encode_crc(original_len=64 << 20), 5, 3)
        a4 = 14 * S4 * n4
        sub["rs104_encode_plus_14_crc_64MiB_stripes"] = {"GBps_algorithmic": round(a4 * world / ms4 / 1e6, 1), "ms": round(ms4, 3),
                                                        "stripes_per_gpu": n4, "scaling": "strong (1024 stripes = 64 GiB of data in total)", "roofline_frac": round(a4 / ms4 / 1e6 / peak, 4),
                                                        "launches_per_step": None,
                                                        "note": "ONE pass over HBM: 2-CTA-cluster byte-LUT encode with the 14 shard CRCs accumulated in registers (sparse-fold CRC), "
                                                                "then the small finisher kernels (1/R of the data); CRC(original) by 32-bit algebra"}
        sub["rs104_encode_plus_14_crc_64MiB_stripes"]["launches_per_step"] = int(l4) // 5
        ms4e, _ = timed(b4.encode, 5, 3)
        sub["rs104_encode_only"] = {"GBps_algorithmic": round(a4 * world / ms4e / 1e6, 1), "roofline_frac": round(a4 / ms4e / 1e6 / peak, 4)}
        del b4
        torch.cuda.empty_cache()
        sub["e2e_32_streams_256MiB_blocks"] = run_config5(L, oracle, np, C, check, rank)
        # small calls: there is no CPU fallback, so a 28-byte encode also goes to the GPU;
        # this is the price in latency, reported so nobody has to guess it
        from gpu_ec import checksum as _ck, erasure as _er
        small = {}
        for name, fn in (("encode_28B_rs42", lambda: _er.encode(b"Hello, Erasure Coding World!", 4, 2)),
                         ("encode_1MiB_rs63", lambda: _er.encode(src_1m, 6, 3)),
                         ("crc32_64KiB", lambda: _ck.crc32(src_1m[:65536])),
                         ("calculate_checksums_1MiB", lambda: _ck.calculate_checksums(src_1m))):
            src_1m = oracle.fill_synthetic(1 << 20, 7)
            for _ in range(5):
                fn()
            t0 = time.perf_counter()
            for _ in range(50):
                fn()
            small[name] = round((time.perf_counter() - t0) / 50 * 1e6, 1)
        sub["small_call_latency_us"] = dict(small, note="host-pointer API incl. pageable H2D/D2H and Python ctypes overhead")
        # many small WriteBlocks (chunkserver.rs:746-769 + :195 each): one batched call vs one call per block vs the CPU port
        bb = {}
        for size, nb in ((4096, 512), (65536, 512), (1 << 20, 128)):
            blocks = [oracle.fill_synthetic(size, 100 + i) for i in range(nb)]
            _ck.block_checksums_batch(blocks)
            t0 = time.perf_counter()
            for _ in range(3):
                wb, cb = _ck.block_checksums_batch(blocks)
            tb = (time.perf_counter() - t0) / 3
            t0 = time.perf_counter()
            for b_ in blocks[:64]:
                _ck.block_checksums(b_)
            t1 = (time.perf_counter() - t0) / 64
            t0 = time.perf_counter()
            for b_ in blocks[:64]:
                oracle.crc32(b_)
                oracle.crc32_chunks(b_)
            tc = (time.perf_counter() - t0) / 64
            if int(wb[0]) != oracle.crc32(blocks[0]) or not np.array_equal(cb[-1], oracle.crc32_chunks(blocks[-1])):
                raise SystemExit("PARITY FAILURE batched block checksums")
            bb[f"{size}B"] = {"batched_blocks_per_s": round(nb / tb), "single_call_blocks_per_s": round(1 / t1), "cpu_port_1thread_blocks_per_s": round(1 / tc),
                              "blocks_per_batch": nb}
        sub["block_checksums_batch"] = dict(bb, note="gpu_ec_block_checksums_batch: pageable host blocks, one CRC launch + one fold launch + one sync per batch")

    if rank == 0:
        traffic, traffic_src = ncu_traffic("rs_lut8_kernel", alg_bytes_rank)
        # the other BASELINE configs, in a key the driver keeps: one roofline entry per dominant kernel
        others = []
        def other(workload, kernel, gbps_key="GBps_algorithmic", frac_key="roofline_frac", alg=None):
            e = sub.get(workload)
            if e:
                others.append({"workload": workload, "kernel": kernel, "achieved": round(e[gbps_key] / world, 1), "unit": "GB/s",
                               "frac": e[frac_key], "peak": peak, "algorithmic_bytes_per_launch": alg})
        if sub:
            other("rs63_reconstruct_3_erasures_4096x1MiB", "rs_lut8_kernel<6> (decode rows)", alg=9 * gpu_ec.erasure.shard_len(1 << 20, 6) * 4096)
            for chunk in (512, 65536):
                w = f"crc32_ieee_{chunk}B_chunks_1GiB"
                other(w, "crc_segments_kernel (write)", "write_GBps", "roofline_frac_write", (1 << 30) + 4 * ((1 << 30) // chunk))
                other(w, "crc_segments_kernel (verify)", "verify_GBps", "roofline_frac_verify", (1 << 30) + 4 * ((1 << 30) // chunk))
            other("rs63_encode_plus_9_crc_64MiB_blocks", "rs_lut8_crc_kernel<6,384> + finisher", alg=alg_bytes_rank)
            other("rs104_encode_only", "rs_x2_crc_fast_kernel<5,384,false>", alg=a4)
            other("rs104_encode_plus_14_crc_64MiB_stripes", "rs_x2_crc_fast_kernel<5,384,true> + finisher", alg=a4)
        line = {
            "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic (splitmix64 counter stream)",
            "config": make_config(n_local, S),
            "gpu_launches": int(launches),
            "clocks": clocks,
            "e2e": None if e2e_val is None else {
                "value": round(e2e_val, 2), "unit": UNIT, "h2d_bytes_per_step": n_e2e * STRIPE_LEN,
                "d2h_bytes_per_step": n_e2e * M_PAR * S, "ms_per_step": round(e2e_ms, 3), "steps": e2e_steps, "stripes_per_step": n_e2e,
                "api": "gpu_ec_encode_batch (pinned host buffers)", "host_binding": numa, "data_GBps": round(n_e2e * STRIPE_LEN * world / e2e_ms / 1e6, 2),
                # e2e `value` = (data + parity) bytes / time = all bytes that crossed PCIe per second
                "pcie_ceiling_GBps_aggregate": PCIE_CEILING.get(world),
                "frac_of_h2d_ceiling": None if world not in PCIE_CEILING else round(n_e2e * STRIPE_LEN * world / e2e_ms / 1e6 / PCIE_CEILING[world]["h2d"], 3),
                "frac_of_bidirectional_ceiling": None if world not in PCIE_CEILING else round(e2e_val / PCIE_CEILING[world]["both"], 3)},
            "roofline": {"bound": "hbm", "achieved": round(per_gpu, 1), "peak": peak, "unit": "GB/s", "frac": round(per_gpu / peak, 4),
                         "traffic": traffic, "traffic_unit": "bytes per launch", "traffic_source": traffic_src,
                         "kernel": "rs_lut8_kernel<6>", "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": alg_bytes_rank, "others": others},
            "sub": sub,
        }
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_reference(3, warmup, False)
            if sub:
                cpu_o = cpu_others(oracle, np)
                line["cpu_baseline"]["others"] = cpu_o
                for o in others:
                    c = cpu_o.get(o["workload"])
                    if c:
                        o["cpu_port_GBps"] = c.get("cpu_port_GBps")
                        o["cpu_cores"] = c.get("cpu_cores")
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
