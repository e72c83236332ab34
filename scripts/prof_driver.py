#!/usr/bin/env python
"""Short single-GPU driver for ncu captures: two builds of the workload text + one query batch
through the C ABI (no CPU baseline, no oracle).  python scripts/prof_driver.py [workload]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from suffixarray_b200 import _lib as S  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "c2"
text, qbytes, qoff, lower = bench.make_workload(name, 0)
idx = S.Index(32, device=0)
for _ in range(2):
    idx.build_text(text, lowercase=lower, timing=True)
st = idx.stats()
s, e, off = idx.query_ranges_packed(qbytes, qoff, k=1000)
pos, doc = idx.query_hits(int(off[-1]))
print({k: st[k] for k in ("n", "rounds", "radix_passes", "kernel_launches", "refined_elems", "finish_runs", "ms_total")},
      "hits", int(off[-1]), "found", int((s != 0xFFFFFFFF).sum()))
idx.close()
