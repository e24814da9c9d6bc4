"""Per-source-line hot spots of one kernel from an .ncu-rep (needs -lineinfo and --import-source on).

    python scripts/ncu_hotspots.py gpurun_out/prof_fused.ncu-rep > profiles/rNN_ncu_source_hotspots.txt
"""
import csv
import subprocess
import sys

rep = sys.argv[1]
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"],
                     capture_output=True, text=True).stdout
rows = list(csv.reader(txt.splitlines()))
hi = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
h = rows[hi]
ix = {}
for i, n in enumerate(h):
    ix.setdefault(n, i)
cols = ["# Samples", "Instructions Executed", "L1 Wavefronts Shared", "L1 Wavefronts Shared Ideal", "stall_short_sb",
        "stall_barrier", "stall_long_sb", "stall_wait", "stall_selected", "stall_not_selected", "stall_dispatch",
        "stall_math", "stall_mio"]
tot = {c: 0.0 for c in cols}
data = []
for r in rows[hi + 1:]:
    if len(r) < len(h) or r[2] != "-":      # keep the per-source-line summary rows
        continue
    try:
        v = {c: float(r[ix[c]] or 0) for c in cols}
    except ValueError:
        continue
    for c in cols:
        tot[c] += v[c]
    data.append((int(r[0]), r[1].strip()[:90], v))
print(rows[1][1] if len(rows) > 1 and len(rows[1]) > 1 else "")
print("totals:", {k: round(v) for k, v in tot.items()})
print("line samples  inst(M) smem_wf(M) ideal(M) short_sb barrier long_sb   wait selected | source")
for ln, src, v in data:
    if v["# Samples"] > tot["# Samples"] * 0.004 or v["L1 Wavefronts Shared"] - v["L1 Wavefronts Shared Ideal"] > 2e6:
        print(f"{ln:4d} {v['# Samples']:7.0f} {v['Instructions Executed'] / 1e6:8.1f} {v['L1 Wavefronts Shared'] / 1e6:10.1f} "
              f"{v['L1 Wavefronts Shared Ideal'] / 1e6:8.1f} {v['stall_short_sb']:8.0f} {v['stall_barrier']:7.0f} "
              f"{v['stall_long_sb']:7.0f} {v['stall_wait']:6.0f} {v['stall_selected']:8.0f} | {src}")
