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


def cpu_reference(steps, warmup, full_line, n_gpus=1):
    """The reference arm / cpu_baseline: C oracle (port) on all host cores, bounded sample."""
    import numpy as np
    from oracle import oracle
    threads = os.cpu_count() or 1
    n_stripes = max(4, min(threads, 16))
    data = oracle.fill_synthetic(STRIPE_LEN * n_stripes, SEED)
    S = oracle.shard_len(STRIPE_LEN, K_DATA)
    alg = (K_DATA + M_PAR) * S * n_stripes
    for _ in range(min(warmup, 1)):
        oracle.bench_encode(data[:STRIPE_LEN * min(n_stripes, threads)], STRIPE_LEN, min(n_stripes, threads), K_DATA, M_PAR, threads)
    ts = [oracle.bench_encode(data, STRIPE_LEN, n_stripes, K_DATA, M_PAR, threads) for _ in range(max(1, steps))]
    t = sum(ts) / len(ts)
    val = alg / t / 1e9
    t1 = oracle.bench_encode(data, STRIPE_LEN, 1, K_DATA, M_PAR, 1)  # what ONE reference call gets: it is single-threaded
    base = {"value": round(val, 3), "unit": UNIT, "cores": threads, "kind": "port",
            "single_thread_value": round((K_DATA + M_PAR) * S / t1 / 1e9, 3),
            "sample": f"{n_stripes} x 64 MiB RS(6,3) stripes, one stripe per thread, {len(ts)} pass(es); "
                      "C restatement of reed-solomon-erasure 6.0.0 scalar table path incl. erasure.rs:17-20 copies"}
    if not full_line:
        return base
    return {"metric": METRIC, "value": round(val, 3), "unit": UNIT, "impl": "reference", "n_gpus": n_gpus, "steps": len(ts),
            "warmup": min(warmup, 1), "ms_per_step": round(t * 1e3, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": "rs63_encode_64MiB_blocks", "k": K_DATA, "m": M_PAR, "stripe_len": STRIPE_LEN,
                       "stripes_per_step": n_stripes},
            "cpu_baseline": base,
            "e2e": {"value": round(val, 3), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}


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

    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    e2e_steps = 3
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

    return e2e_val, e2e_ms


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
    want = oracle.encode(srcs[0][1], k, m)  # parity gate on one block (thread 0 used source 0)
    got = pars[0][1].reshape(m, S)
    ok = np.array_equal(got, want[k:]) and crcs[0].tolist() == [zlib.crc32(x.tobytes()) for x in want]
    ok = ok and fulls[0].value == zlib.crc32(srcs[0][1].tobytes())
    if not ok:
        raise SystemExit("PARITY FAILURE config 5")
    for p, _ in srcs + pars:
        L.gpu_ec_free_pinned(p)
    total = n_thr * iters * blk
    return {"threads": n_thr, "block_bytes": blk, "blocks_per_thread": iters,
            "write_data_GBps": round(total / tw / 1e9, 2), "read_verify_data_GBps": round(total / tr / 1e9, 2),
            "note": "per GPU, wall clock, PCIe-bound: write moves 1.0x data H2D + 0.5x D2H, read moves 1.0x H2D"}


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
            print(json.dumps(cpu_reference(args.steps, args.warmup, True, args.gpus)), flush=True)
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
    batch.fill_synthetic(SEED + rank)
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
    e2e_val = e2e_ms = None
    if not args.no_e2e:
        e2e_val, e2e_ms = run_e2e(L, oracle, np, C, check, barrier, partition, dist, world, rank, S, n_e2e)
    # ---- secondary configs (device-resident), each with its own roofline fraction
    sub = {}
    if not args.no_sub:
        del batch
        torch.cuda.empty_cache()
        # config 3: 4096 stripes of 1 MiB, RS(6,3), shards {1,4,7} erased -> one pass, reads k writes e
        S3 = gpu_ec.erasure.shard_len(1 << 20, 6)
        b3 = device.StripeBatch(4096, 6, 3, S3)
        b3.fill_synthetic(0xEC000003 + rank)
        b3.encode()
        ref3 = b3.buf[:8].clone()
        b3.buf[:, [1, 4, 7]] = 0
        present = [1, 0, 1, 1, 0, 1, 1, 0, 1]
        b3.reconstruct(present)
        torch.cuda.synchronize()
        if not torch.equal(b3.buf[:8, :, :S3], ref3[:, :, :S3]):
            raise SystemExit("PARITY FAILURE reconstruct")
        ms3, _ = timed(lambda: b3.reconstruct(present), 10, 3)
        a3 = 9 * S3 * 4096
        sub["rs63_reconstruct_3_erasures_4096x1MiB"] = {"GBps_algorithmic": round(a3 * world / ms3 / 1e6, 1), "ms": round(ms3, 4),
                                                       "roofline_frac": round(a3 / ms3 / 1e6 / peak, 4)}
        del b3, ref3
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
        if world == 1 and not args.no_cpu:  # CPU side of config 2 on a bounded sample (crc32fast's PCLMULQDQ path restated)
            threads = os.cpu_count() or 1
            samp = t2[:256 << 20].cpu().numpy()
            for chunk in (512, 65536):
                tc = oracle.bench_crc_chunks(samp, chunk, 0, threads)
                t1c = oracle.bench_crc_chunks(samp[:32 << 20], chunk, 0, 1)
                sub[f"crc32_ieee_{chunk}B_chunks_1GiB"]["cpu_port_GBps"] = round(samp.size / tc / 1e9, 2)
                sub[f"crc32_ieee_{chunk}B_chunks_1GiB"]["cpu_port_1thread_GBps"] = round((32 << 20) / t1c / 1e9, 2)
                sub[f"crc32_ieee_{chunk}B_chunks_1GiB"]["cpu_cores"] = threads
            del samp
        host_sample = t2[:1 << 20].cpu().numpy()
        chk = device.u32(device.crc32_chunks(t2[:1 << 20], 512))
        if not np.array_equal(chk, oracle.crc32_chunks(host_sample, 512)):
            raise SystemExit("PARITY FAILURE crc chunks")
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
        ms4, _ = timed(lambda: b4.encode_crc(original_len=64 << 20), 5, 3)
        a4 = 14 * S4 * n4
        sub["rs104_encode_plus_14_crc_64MiB_stripes"] = {"GBps_algorithmic": round(a4 * world / ms4 / 1e6, 1), "ms": round(ms4, 3),
                                                        "stripes_per_gpu": n4, "scaling": "strong (1024 stripes = 64 GiB of data in total)", "roofline_frac": round(a4 / ms4 / 1e6 / peak, 4),
                                                        "note": "encode pass, then segment-CRC + fold kernels over all 14 shards; CRC(original) by 32-bit algebra; not one fused kernel"}
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

    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 2), "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warmup,
            "ms_per_step": round(ms, 4), "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u8",
            "data": "synthetic (splitmix64 counter stream)",
            "config": {"workload": "rs63_encode_64MiB_blocks", "k": K_DATA, "m": M_PAR, "stripe_len": STRIPE_LEN,
                       "shard_len": S, "stripes_per_gpu": n_local, "partition": "stripe i -> gpu i mod G, no collective",
                       "l2": "inputs (6.4 GB/GPU) exceed the 126 MB L2; no flush needed", "bytes_counted": "(k+m)*shard_len per stripe"},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "e2e": None if e2e_val is None else {
                "value": round(e2e_val, 2), "unit": UNIT, "h2d_bytes_per_step": n_e2e * STRIPE_LEN,
                "d2h_bytes_per_step": n_e2e * M_PAR * S, "ms_per_step": round(e2e_ms, 3), "stripes_per_step": n_e2e,
                "api": "gpu_ec_encode_batch (pinned host buffers)", "host_binding": numa, "data_GBps": round(n_e2e * STRIPE_LEN * world / e2e_ms / 1e6, 2)},
            "roofline": {"bound": "hbm", "achieved": round(per_gpu, 1), "peak": peak, "unit": "GB/s", "frac": round(per_gpu / peak, 4),
                         # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of this shape (64 stripes) from the
                         # committed ncu --set full capture profiles/r01_final_ncu_full_summary.txt (4.295 GB + 2.117 GB)
                         "traffic": 6412392000 if n_local == 64 else None, "traffic_unit": "bytes per launch",
                         "kernel": "rs_lut8_kernel<6>", "peak_source": peak_src,
                         "algorithmic_bytes_per_launch": alg_bytes_rank},
            "sub": sub,
        }
        if world == 1 and not args.no_cpu:
            line["cpu_baseline"] = cpu_reference(1, 0, False)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
