#!/usr/bin/env python
"""Hot SASS instructions of the first kernel in an .ncu-rep (source page, needs -lineinfo):
    python profiles/ncu_hot.py rep.ncu-rep [top_n] [launch_index | kernel-name substring]
Prints the instructions with the most stall samples and the dominant stall reason of each."""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 25
which = sys.argv[3] if len(sys.argv) > 3 else "0"
txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
blocks, cur = [], None
for row in csv.reader(io.StringIO(txt)):
    if row and row[0] == "Kernel Name":
        cur = {"name": row[1], "hdr": None, "rows": []}
        blocks.append(cur)
    elif cur is not None and row and row[0] == "Address":
        cur["hdr"] = row
    elif cur is not None and cur["hdr"] and len(row) == len(cur["hdr"]):
        cur["rows"].append(row)
blocks = [b for b in blocks if b["rows"]]
b = blocks[int(which)] if which.isdigit() else [b for b in blocks if which in b["name"]][0]
h = {k: i for i, k in enumerate(b["hdr"])}
stall_cols = [k for k in b["hdr"] if k.startswith("stall_") and "Not Issued" not in k]
tot = sum(float(r[h["# Samples"]] or 0) for r in b["rows"])
print(b["name"][:100], "total samples", tot, "instructions", len(b["rows"]))
agg = {k: sum(float(r[h[k]] or 0) for r in b["rows"]) for k in stall_cols}
print("stall totals:", ", ".join(f"{k[6:]} {v / tot * 100:.1f}%" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
rows = sorted(enumerate(b["rows"]), key=lambda ir: -float(ir[1][h["# Samples"]] or 0))[:top]
for i, r in sorted(rows):
    s = float(r[h["# Samples"]] or 0)
    reasons = sorted(((float(r[h[k]] or 0), k[6:]) for k in stall_cols), reverse=True)[:2]
    print(f"{i:5d} {s / tot * 100:5.1f}%  {r[h['Source']][:70]:70s} " + " ".join(f"{n}:{v:.0f}" for v, n in reasons))
