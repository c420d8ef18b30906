#!/usr/bin/env python3
"""Summarise an .ncu-rep (run here, no GPU needed) into the text that gets committed under profiles/.

usage: python tools/ncu_summary.py gpurun_out/x.ncu-rep profiles/r02/x_ncu_summary.txt [--mix TILES]
  per captured launch: duration, DRAM bytes, pipe utilisation, issue rate, stall reasons per issue;
  --mix N additionally prints the executed-instruction mix per N (e.g. warp-tiles) from the source page.
"""
import csv
import io
import re
import subprocess
import sys
from collections import Counter

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__cluster_size",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__warps_active.avg.pct_of_peak_sustained_active"]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    mix = int(sys.argv[sys.argv.index("--mix") + 1]) if "--mix" in sys.argv else 0
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    H, U = rows[0], rows[1]
    lines = [f"# {rep} — extracted with `ncu -i ... --page raw --csv` (tools/ncu_summary.py)"]
    seen = set()
    for r in rows[2:]:
        d = dict(zip(H, r))
        u = dict(zip(H, U))
        key = (d.get("Kernel Name"), d.get("Grid Size"), d.get("Block Size"))
        if key in seen:   # one entry per kernel and launch shape (repeat launches differ by a fraction of a percent)
            continue
        seen.add(key)
        lines.append("")
        lines.append(f"== {d.get('Kernel Name', '?')}  grid {d.get('Grid Size', '?')} block {d.get('Block Size', '?')}")
        for k in KEYS:
            if k in d and d[k] != "":
                lines.append(f"{k} = {d[k]} {u.get(k, '')}")
        for k in H:
            if "issue_stalled" in k and k.endswith("per_issue_active.ratio"):
                try:
                    if float(d[k]) >= 0.05:
                        lines.append(f"{k} = {float(d[k]):.3f}")
                except ValueError:
                    pass
    if mix:
        src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(src)))
        hdr = next(i for i, r in enumerate(rows) if "Source" in r and "Instructions Executed" in r)
        si, ei = rows[hdr].index("Source"), rows[hdr].index("Instructions Executed")
        c, tot = Counter(), 0
        for r in rows[hdr + 1:]:
            try:
                n = int(r[ei])
            except (ValueError, IndexError):
                continue
            t = re.sub(r"^@!?U?P\d\s+", "", r[si].strip())
            op = t.split()[0].split(".")[0] + (".HI" if "IMAD.HI" in t else "")
            c[op] += n
            tot += n
        lines.append("")
        lines.append(f"executed warp instructions per unit ({mix} units, first kernel of the source page): {tot / mix:.1f}")
        lines.append("  " + ", ".join(f"{op} {n / mix:.1f}" for op, n in c.most_common(20)))
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
