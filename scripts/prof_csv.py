#!/usr/bin/env python
"""Short single-GPU driver for ncu launch lists of the CSV path: one synthetic file (--csv-mb of
`name` column), sa_build_csv twice, sa_build_csv_columns over three columns once, one query batch
with hits (file offsets + rows).  python scripts/prof_csv.py [--csv-mb 256]"""
import argparse
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from scripts.sweeps import write_csv  # noqa: E402
from suffixarray_b200 import _lib as S  # noqa: E402
from suffixarray_b200 import workloads as W  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--csv-mb", type=int, default=256)
args = ap.parse_args()
path = os.path.join(tempfile.gettempdir(), "sa_prof.csv")
if not os.path.exists(path):
    write_csv(path, args.csv_mb)
idx = S.Index(32, device=0)
for _ in range(2):
    idx.build_csv(path, "name", timing=True)
st = idx.stats()
qbytes, qoff = W.substring_queries(idx.text(), 200_000, seed=3002)
s, e, off = idx.query_ranges_packed(qbytes, qoff, k=1000)
pos, row = idx.query_hits(int(off[-1]))
idxs = [idx, S.Index(32, device=0), S.Index(32, device=0)]
S.Index.build_csv_columns(idxs, path, ["name", "city", "score"], timing=True)
print({k: st[k] for k in ("n", "ms_total", "ms_h2d", "ms_csv_scan", "ms_remap", "kernel_launches")}, "hits", int(off[-1]),
      "rows max", int(row.max()))
for ix in idxs:
    ix.close()
