#!/usr/bin/env python
"""BASELINE config 3 over N GPUs, run under torchrun (one process per GPU):
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/check_dist_csv.py [--csv-mb 256]
Rank 0 writes a synthetic CSV; every rank then runs its part of ONE index over the `name` column
(sa_shard_begin_csv: per-rank device scan of the file, sample sort with peer stores, finish +
replicate, remap to file offsets), timed from open() to a queryable index on every rank.  The array
of every rank is compared with a single-GPU sa_build_csv of the same file (digest), as are the
quoted bits, and a query batch is answered on every rank."""
import argparse
import hashlib
import json
import os
import sys
import tempfile
import time

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from suffixarray_b200 import _lib as S  # noqa: E402
from suffixarray_b200 import sharded as SH  # noqa: E402
from suffixarray_b200 import workloads as W  # noqa: E402
from scripts.sweeps import write_csv  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--csv-mb", type=int, default=256)
ap.add_argument("--queries", type=int, default=200_000)
ap.add_argument("--out", default="")
args = ap.parse_args()

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
L = 32
path = os.path.join(tempfile.gettempdir(), "sa_dist_c3.csv")
rows = fbytes = 0
if rank == 0:
    rows, fbytes = write_csv(path, args.csv_mb)
dist.barrier()


def digest(a: np.ndarray) -> bytes:
    return hashlib.blake2b(np.ascontiguousarray(a).tobytes(), digest_size=8).digest()


builder = None
walls = []
for rep in range(3):
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    builder, phases = SH.build_distributed(None, L, local, timing=True, builder=builder, csv=(path, "name"))
    torch.cuda.synchronize()
    dist.barrier()
    walls.append((time.perf_counter() - t0) * 1e3)
n = builder.idx.n
mine = builder.idx.sa()
dg = torch.tensor(list(digest(mine) + digest(builder.idx.quoted_bits())), dtype=torch.int32, device="cuda")
all_dg = [torch.zeros_like(dg) for _ in range(world)]
dist.all_gather(all_dg, dg)
replicas_equal = all(bool(torch.equal(all_dg[0], x)) for x in all_dg)
# every rank answers its own share of a query batch
text = builder.idx.text()
qbytes, qoff = W.substring_queries(text, args.queries, seed=3002 + rank)
s, e, off = builder.idx.query_ranges_packed(qbytes, qoff, k=1000)
pos, row = builder.idx.query_hits(int(off[-1]))
qms = builder.idx.stats()["ms_query"]
report = None
if rank == 0:
    single = S.Index(L, device=local)
    single_runs = []
    for _ in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        single.build_csv(path, "name", timing=True)
        single_runs.append((time.perf_counter() - t0) * 1e3)
    single_ms = min(single_runs[1:])
    same = bool(np.array_equal(single.sa(), mine)) and bool(np.array_equal(single.quoted_bits(), builder.idx.quoted_bits()))
    s1, e1, off1 = single.query_ranges_packed(qbytes, qoff, k=1000)
    pos1, row1 = single.query_hits(int(off1[-1]))
    qsame = bool(np.array_equal(s, s1) and np.array_equal(e, e1) and np.array_equal(pos, pos1) and np.array_equal(row, row1))
    report = {"world": world, "csv_mb": round(fbytes / 1e6, 1), "rows": rows, "column_chars": int(n),
              "e2e_ms_runs": [round(w, 1) for w in walls], "e2e_ms": round(min(walls[1:]), 1),
              "mchars_per_s_e2e": round(n / min(walls[1:]) / 1e3, 1),
              "phases_ms_last_rank0": {k: round(v, 2) for k, v in phases.items()},
              "single_gpu_e2e_ms": round(single_ms, 1), "single_gpu_device_build_ms": round(single.stats()["ms_total"], 2),
              "array_and_quoted_bits_equal_single_gpu": same, "replicas_equal": replicas_equal,
              "queries_equal_single_gpu": qsame, "range_queries_per_s_device_per_rank": round(args.queries / qms * 1e3, 0),
              "row_ids_in_range": bool(int(row.max()) <= rows)}
    single.close()
    os.remove(path)
    print(json.dumps(report), flush=True)
    if args.out:
        with open(args.out, "w") as f:
            f.write(json.dumps(report) + "\n")
builder.close()
dist.barrier()
dist.destroy_process_group()
ok = report is None or (report["array_and_quoted_bits_equal_single_gpu"] and report["replicas_equal"] and report["queries_equal_single_gpu"])
sys.exit(0 if ok else 1)
