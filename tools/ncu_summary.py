#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU): one table row per profiled kernel instance.
    python tools/ncu_summary.py gpurun_out/r2/evidence.ncu-rep > profiles/r2/ncu_evidence.md
    python tools/ncu_summary.py gpurun_out/r2/evidence_raw.csv  > profiles/r2/ncu_evidence.md"""
import csv, io, subprocess, sys, collections

rep = sys.argv[1]
if rep.endswith(".csv"):  # already exported on the GPU box (`ncu -i rep --page raw --csv`)
    raw = open(rep).read()
else:
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr = rows[0]
col = {h: i for i, h in enumerate(hdr)}
want = [
    ("gpu__time_duration.sum", "time us", 1e-3),
    ("dram__bytes_read.sum", "DRAM rd MB", 1e-6),
    ("dram__bytes_write.sum", "DRAM wr MB", 1e-6),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %", 1),
    ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor %", 1),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %", 1),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ %", 1),
    ("launch__registers_per_thread", "regs", 1),
    ("launch__grid_size", "grid", 1),
    ("launch__block_size", "block", 1),
    ("launch__cluster_size", "cluster", 1),
]
have = [(m, n, s) for m, n, s in want if m in col]
units = rows[1] if len(rows) > 1 else []
print("| kernel | " + " | ".join(n for _, n, _ in have) + " |")
print("|---|" + "---:|" * len(have))
seen = collections.Counter()
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    name = r[col["Kernel Name"]].split("(")[0].split("::")[-1]
    seen[name] += 1
    if seen[name] > 2:
        continue
    cells = []
    for m, n, s in have:
        v = r[col[m]].replace(",", "")
        u = units[col[m]] if units else ""
        try:
            f = float(v)
            if m == "gpu__time_duration.sum":
                f = f * ({"ns": 1e-3, "us": 1.0, "ms": 1e3, "s": 1e6}.get(u, 1e-3))
            elif m.startswith("dram__bytes"):
                f = f * ({"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1e-6))
            cells.append(f"{f:.2f}" if f < 1000 else f"{f:.0f}")
        except ValueError:
            cells.append(v)
    print(f"| `{name}` | " + " | ".join(cells) + " |")
