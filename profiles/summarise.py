#!/usr/bin/env python
"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel totals.

    python profiles/summarise.py gpurun_out/launches.csv [skip_launches] > profiles/rNN_launches.md

`skip_launches`: launches to ignore at the front (the warm-up step).  Times under ncu are
cold-cache and serialised, so only each kernel's SHARE of the step is meaningful.
"""
import csv
import re
import sys
from collections import OrderedDict


def main():
    path = sys.argv[1]
    skip = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    rows = []
    with open(path, newline="") as f:
        lines = [ln for ln in f if not ln.startswith("==")]
    for r in csv.DictReader(lines):
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        name = re.sub(r"\(.*", "", r["Kernel Name"]).replace("sa::", "")
        val = float(r["Metric Value"].replace(",", ""))
        unit = r["Metric Unit"]
        ns = val * {"ns": 1, "us": 1e3, "ms": 1e6, "s": 1e9}.get(unit, 1)
        rows.append((int(r["ID"]), name, ns, r["Grid Size"], r["Block Size"]))
    rows = rows[skip:]
    agg = OrderedDict()
    for _, name, ns, grid, block in rows:
        a = agg.setdefault(name, [0, 0.0, 0.0])
        a[0] += 1
        a[1] += ns
        a[2] = max(a[2], ns)
    total = sum(a[1] for a in agg.values())
    print(f"launches: {len(rows)} (skipped {skip}), total kernel time {total / 1e6:.3f} ms\n")
    print("| kernel | launches | total ms | share | avg us | max us |")
    print("|---|---:|---:|---:|---:|---:|")
    for name, (c, t, mx) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"| `{name}` | {c} | {t / 1e6:.3f} | {100 * t / total:.1f}% | {t / c / 1e3:.1f} | {mx / 1e3:.1f} |")


if __name__ == "__main__":
    main()
