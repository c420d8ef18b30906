#!/usr/bin/env python3
"""Build tools/ref_harness with cargo (needs a Rust toolchain AND the two crates in a local registry
cache: there is no network) and regenerate tests/golden/ref_golden.json from the REAL crates, on the
inputs tests/golden/golden.json already holds. Exits 3 with one line when cargo is missing — which is
the case in this image and on the GPU boxes (probe: profiles/r02/r02_toolchain_probe.txt)."""
import json
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))

if not shutil.which("cargo"):
    print("cargo not found: the reference's crates cannot be built here; parity stays pinned to external vectors only")
    sys.exit(3)
subprocess.check_call(["cargo", "build", "--release", "--offline"], cwd=HERE)
exe = os.path.join(HERE, "target", "release", "ref_harness")
golden = json.load(open(os.path.join(ROOT, "tests", "golden", "golden.json")))
out = {"generator": "tools/ref_harness (reed-solomon-erasure 6.0.0, crc32fast 1.5.0)", "encode_cases": []}
for c in golden["encode_cases"]:
    lines = subprocess.check_output([exe, "encode", str(c["k"]), str(c["m"]), c["data"]], text=True).split("\n")
    rows = [ln.split() for ln in lines if ln]
    out["encode_cases"].append({"k": c["k"], "m": c["m"], "data": c["data"], "shards": [r[1] for r in rows],
                                "shard_crc_ieee": [int(r[2]) for r in rows]})
json.dump(out, open(os.path.join(ROOT, "tests", "golden", "ref_golden.json"), "w"), indent=1)
print("wrote tests/golden/ref_golden.json")
