#!/usr/bin/env python
"""Multi-GPU parity check, run under torchrun (one process per GPU):
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/check_dist.py
Builds ONE suffix array over N GPUs (sample sort + NCCL all-to-all) for several texts and compares
the adopted array on every rank with the oracle, bit for bit."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import pyoracle as O  # noqa: E402
from suffixarray_b200 import sharded as SH  # noqa: E402
from suffixarray_b200 import workloads as W  # noqa: E402

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
rng = np.random.default_rng(7)
cases = [
    ("english 3M L32", W.english_documents(3_000_000, seed=61), 32, False),
    ("names 100k L32 upper", W.company_names_joined(100_000, seed=62, upper=True)[0], 32, True),
    ("sigma2 200k L16", (rng.integers(0, 2, 200_000) + 97).astype(np.uint8), 16, False),
    ("bytes 100k L32", rng.integers(0, 256, 100_000).astype(np.uint8), 32, False),
    ("tiny 50", (rng.integers(0, 3, 50) + 97).astype(np.uint8), 32, False),
]
ok = True
builder = None
for name, text, L, lower in [c for c in cases for _ in (0, 1)]:
    exchange = "nccl" if getattr(builder, "_last", None) == (name, "p2p") else "p2p"
    if builder is not None and builder.L != L:
        builder.close()
        builder = None
    replicate = "fused" if exchange == "p2p" else "allgather"
    builder, phases = SH.build_distributed(text, L, local, lowercase=lower, builder=builder, exchange=exchange,
                                           replicate=replicate)
    builder._last = (name, exchange)
    got = builder.idx.sa()
    want_text = O.oracle_lowercase(text) if lower else text
    want = O.oracle_build(want_text, L)
    same = bool(np.array_equal(got, want))
    flags = torch.tensor([int(same)], device="cuda")
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"{name} [{exchange}/{replicate}]: n={text.size} world={world} shard m={builder.m} -> {'OK' if flags.item() else 'MISMATCH'}  "
              + " ".join(f"{k}={v:.1f}ms" for k, v in phases.items()), flush=True)
    ok = ok and bool(flags.item())
builder.close()
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
