#!/usr/bin/env python3
"""Static SASS instruction mix per kernel of libgpu_ec.so (cuobjdump -sass), whole kernel and its largest
loop body (the streaming loop). `make -C rust-hadoop-generated-by-llm_b200/csrc sass` writes profiles/sass_static_mix.txt.
Static counts include cold paths (edge tiles, prologue); the executed mix per tile is in the ncu summaries."""
import re
import subprocess
import sys
from collections import Counter

so = sys.argv[1]
txt = subprocess.run(["cuobjdump", "-sass", so], capture_output=True, text=True).stdout
out = [f"# static SASS mix of {so.split('/')[-1]} (sm_100a), tools/sass_mix.py"]
for f in re.split(r"\n\s*Function : ", txt)[1:]:
    name = subprocess.run(["c++filt", f.split("\n")[0].strip()], capture_output=True, text=True).stdout.strip()
    if "gpu_ec::" not in name:
        continue
    ins = []
    for l in f.split("\n"):
        m = re.match(r"\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);", l)
        if m:
            ins.append((int(m.group(1), 16), m.group(2).strip()))
    def mix(body):
        c = Counter(re.sub(r"^@!?U?P\d\s+", "", t).split()[0].split(".")[0] for _, t in body)
        return ", ".join(f"{k} {v}" for k, v in c.most_common(12))
    best = []
    for a, t in ins:
        m = re.search(r"BRA(?:\.U)?\s+(?:!?U?P\d,\s*)?0x([0-9a-f]+)", t)
        if m and int(m.group(1), 16) < a:
            body = [x for x in ins if int(m.group(1), 16) <= x[0] <= a]
            if len(body) > len(best):
                best = body
    special = sorted({re.sub(r"^@!?U?P\d\s+", "", t).split()[0].split(".")[0] for _, t in ins} &
                     {"LDGSTS", "STAS", "SYNCS", "UTMALDG", "UTMASTG", "UBLKCP", "UCGABAR_ARV", "SHFL", "REDUX", "LDS", "STS"})
    out.append(f"\n== {name.split('(')[0]}\n   whole kernel: {len(ins)} instr: {mix(ins)}\n   largest loop: {len(best)} instr: {mix(best)}\n   notable: {' '.join(special)}")
print("\n".join(out))
