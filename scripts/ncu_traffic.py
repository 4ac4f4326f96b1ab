#!/usr/bin/env python
"""Extract per-launch DRAM traffic and duration of the level-0 SpMV kernel from an `ncu --set full`
report of scripts/prof_ops.py and write the committed summary that bench.py quotes as roofline.traffic.
usage: ncu_traffic.py gpurun_out/prof_rows.ncu-rep profiles/r01_ncu_traffic.json [rows nnz value_format]"""
import csv
import hashlib
import io
import json
import os
import subprocess
import sys

rep, out = sys.argv[1], sys.argv[2]
rows_n = int(sys.argv[3]) if len(sys.argv) > 3 else 256 ** 3
nnz = int(sys.argv[4]) if len(sys.argv) > 4 else 117047296
vfmt = int(sys.argv[5]) if len(sys.argv) > 5 else 0
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
col = {h: i for i, h in enumerate(hdr)}


def val(r, name):
    v = float(r[col[name]])
    u = units[col[name]]
    scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "us": 1e-6, "ms": 1e-3, "ns": 1e-9, "s": 1.0}.get(u, 1.0)
    return v * scale


launches = []
for r in rows[2:]:
    name = r[col["Kernel Name"]]
    launches.append({
        "kernel": name[:120],
        "duration_us": val(r, "gpu__time_duration.sum") * 1e6,
        "dram_read_bytes": val(r, "dram__bytes_read.sum"),
        "dram_write_bytes": val(r, "dram__bytes_write.sum"),
        "registers": int(float(r[col["launch__registers_per_thread"]])),
        "warps_active_pct": float(r[col["sm__warps_active.avg.pct_of_peak_sustained_active"]]),
        "dram_cycles_active_pct": float(r[col["dram__cycles_active.avg.pct_of_peak_sustained_elapsed"]]),
    })
spmv = [l for l in launches if "EpiMatvec>" in l["kernel"] or "EpiMatvec," in l["kernel"] or "9EpiMatvec" in l["kernel"]]
first = (spmv or launches)[0]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
summary = {
    # bench.py quotes the traffic only while the kernel source is the one that was profiled
    "spmv_cuh_sha16": hashlib.sha256(open(os.path.join(ROOT, "hypre_b200", "csrc", "spmv.cuh"), "rb").read()).hexdigest()[:16],
    "commit": subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip(),
    "source": "ncu --set full --clock-control none -k regex:rows_kernel -s 3 -c 4 python scripts/prof_ops.py --reps 2",
    "matrix": {"rows": rows_n, "nnz": nnz, "value_format": vfmt, "what": "7-pt Laplacian 256^3, level-0 diag block, thread-per-row layout"},
    "spmv_level0": {"traffic_bytes_per_launch": first["dram_read_bytes"] + first["dram_write_bytes"],
                    "algorithmic_bytes_per_launch": 12.0 * nnz + 4.0 * (rows_n + 1) + 16.0 * rows_n, **first},
    "launches": launches,
}
json.dump(summary, open(out, "w"), indent=1)
print(json.dumps(summary["spmv_level0"]))
