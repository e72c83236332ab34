#!/usr/bin/env python
"""Reader threads of the CSV file ring (SA_CSV_READERS, read once per process): file read + H2D time
of sa_build_csv on one synthetic file.  python scripts/csv_readers_sweep.py  (spawns one child per setting)"""
import json
import os
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
CHILD = r"""
import sys, json, time
sys.path.insert(0, sys.argv[1])
from suffixarray_b200 import _lib as S
import torch
idx = S.Index(32, device=0)
out = []
for _ in range(4):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    idx.build_csv(sys.argv[2], "name", timing=True)
    torch.cuda.synchronize(); w = (time.perf_counter() - t0) * 1e3
    st = idx.stats(); out.append((round(w, 1), round(st["ms_h2d"], 1)))
print("RES" + json.dumps(out))
"""
if __name__ == "__main__":
    from scripts.sweeps import write_csv
    path = os.path.join(tempfile.gettempdir(), "sa_readers.csv")
    rows, size = write_csv(path, int(sys.argv[1]) if len(sys.argv) > 1 else 512)
    res = {"file_mb": round(size / 1e6, 1)}
    for r in (2, 4, 8, 12, 16):
        p = subprocess.run([sys.executable, "-c", CHILD, ROOT, path], env=dict(os.environ, SA_CSV_READERS=str(r)),
                           capture_output=True, text=True)
        line = [ln for ln in p.stdout.splitlines() if ln.startswith("RES")]
        res[f"readers_{r}"] = json.loads(line[0][3:]) if line else p.stderr[-300:]
    os.remove(path)
    print(json.dumps(res))
