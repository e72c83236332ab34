#!/usr/bin/env python
"""Read an .ncu-rep (ncu -i … --page raw --csv) and print, per captured launch group (kernel name),
the metrics the roofline needs plus the warp-stall breakdown.  Usage:
    python profiles/ncu_read.py gpurun_out/prof.ncu-rep [--json out.json]
"""
import csv
import io
import json
import re
import subprocess
import sys
from collections import OrderedDict

KEYS = OrderedDict([
    ("gpu__time_duration.sum", "time"),
    ("dram__bytes_read.sum", "dram_rd"),
    ("dram__bytes_write.sum", "dram_wr"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ_pct"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_pct"),
    ("launch__registers_per_thread", "regs"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__occupancy_limit_registers", "lim_regs"),
    ("launch__occupancy_limit_shared_mem", "lim_smem"),
    ("lts__t_sector_hit_rate.pct", "l2_hit"),
    ("l1tex__t_sector_hit_rate.pct", "l1_hit"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_conflicts"),
])
TO_BYTES = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
TO_US = {"ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}


def main():
    rep = sys.argv[1]
    txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(txt)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    col = {h: i for i, h in enumerate(hdr)}
    stall_cols = [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
    groups = OrderedDict()
    for r in data:
        name = re.sub(r"\(.*", "", r[col["Kernel Name"]]).replace("sa::", "").replace("void ", "")
        groups.setdefault(name, []).append(r)
    out = {}
    for name, rs in groups.items():
        def num(h, r):
            v = r[col[h]].replace(",", "")
            try:
                return float(v)
            except ValueError:
                return float("nan")
        rec = {"launches": len(rs)}
        for h, short in KEYS.items():
            if h not in col:
                continue
            vals = [num(h, r) for r in rs]
            u = units[col[h]]
            if u in TO_BYTES:
                vals = [v * TO_BYTES[u] for v in vals]
            if short == "time":
                vals = [v * TO_US[u] for v in vals]
            rec[short] = sum(vals) / len(vals)
        rec["dram_bytes"] = rec.get("dram_rd", 0) + rec.get("dram_wr", 0)
        rec["dram_gbs"] = rec["dram_bytes"] / (rec["time"] * 1e-6) / 1e9 if rec.get("time") else 0
        stalls = {h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]:
                  sum(num(h, r) for r in rs) / len(rs) for h in stall_cols}
        rec["stalls"] = dict(sorted(stalls.items(), key=lambda kv: -kv[1])[:6])
        out[name] = rec
        print(f"## {name}  ({len(rs)} launches)")
        print(f"  time {rec['time']:.1f} us | DRAM {rec['dram_bytes'] / 1e6:.1f} MB/launch = {rec['dram_gbs']:.0f} GB/s "
              f"({rec.get('dram_pct', 0):.1f}% of ncu peak) | L2 hit {rec.get('l2_hit', 0):.1f}% L1 hit {rec.get('l1_hit', 0):.1f}%")
        print(f"  grid {rec.get('grid', 0):.0f} x {rec.get('block', 0):.0f} | regs {rec.get('regs', 0):.0f} | blocks/SM limit regs "
              f"{rec.get('lim_regs', 0):.0f} smem {rec.get('lim_smem', 0):.0f} | warps active {rec.get('occ_pct', 0):.1f}% | "
              f"issue active {rec.get('issue_pct', 0):.1f}% | SM {rec.get('sm_pct', 0):.1f}% | smem conflicts {rec.get('smem_conflicts', 0):.3g}")
        print("  stalls (warps per issue): " + ", ".join(f"{k} {v:.2f}" for k, v in rec["stalls"].items()))
    if "--json" in sys.argv:
        with open(sys.argv[sys.argv.index("--json") + 1], "w") as f:
            json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
