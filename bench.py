#!/usr/bin/env python
"""bench.py -- SA build Mchars/s + batched substring queries/s (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload c2|c1|c3|c5|small]

A "step" is one pass of the hot path over one batch of synthetic input: build the truncated
suffix array of the workload text (max_suffix_length = 32) and answer the workload's query batch
(ranges + hit lists).  Default workload = BASELINE.json configs[1]: a 100 M-char synthetic
English document list + a 1 M-query batch.

* `value`   : build Mchars/s with the text already resident in HBM (device-timed, CUDA events on
              the library's stream, max over ranks).  `queries_per_s` is the query half of the
              metric on the same footing (query batch resident in HBM).
* `e2e`     : the same two numbers through the C-ABI calls a user of the drop-in makes
              (sa_build_text / sa_query_ranges / sa_query_hits; sa_shard_build at N > 1) with PINNED
              HOST buffers, the host->device and device->host copies inside the timed region.
* `roofline`: the dominant kernel (onesweep radix scatter pass): algorithmic bytes per launch
              (24 B per record moved: read 8+4, write 8+4) / average launch time, against the
              measured HBM copy peak in MEASURED_PEAKS.json.  `roofline_query`: the query kernel
              against SURVEY 8(d)'s B_query (useful and sector bytes).
* `cpu_baseline`: the reference's own engine.c (oracle/_ref/libsa_ref.so, built from
              /root/reference in the authoring container) timed on ALL host cores (thread count set
              explicitly: torchrun exports OMP_NUM_THREADS=1), rank 0, at every N.
* `parity_checked`: the text the CPU arm indexed is built on the GPU too and the arrays, the ranges of
              the whole batch and a sample of hit lists are compared entry by entry.

N > 1 (torchrun, one process per GPU): ONE global index over the concatenation of the N per-rank
texts (weak scaling), built by the library's native driver (sa_shard_build: every rank uploads only
its slice; records, text and array slices move as peer stores over NVLink; small exchanges through a
shared-memory communicator).  torch.distributed / NCCL carries the bench's own barriers and the
max / sum of the timings.  `--driver python` is the round-1 Python loop, `--mode partition` the
reference's own sharding (independent partitions).

`--impl reference` times the reference CPU path on the per-GPU workload of the same config, rank 0 only.
`--workload c3|c5` run BASELINE configs 3 and 5 through scripts/sweeps.py (CPU arm beside every number).
The oracle/ directory is used here ONLY for the cpu_baseline / reference arm and the parity check.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "SA build Mchars/s and batched substring queries/s at 1/2/4/8 B200 vs CPU"
L_DEFAULT = 32
K_HITS = 1000

WORKLOADS = {
    # name: (text chars, queries, description)
    "c2": (100_000_000, 1_000_000, "100M-char synthetic English document list, max_suffix_length=32, 1M-query batch"),
    "c1": (0, 10_000, "1M synthetic company-name strings, max_suffix_length=32, 10k whole-name queries"),
    "small": (4_000_000, 50_000, "4M-char synthetic English (debug size)"),
    # the two below run through scripts/sweeps.py (their own step structure), 1 GPU
    "c3": (0, 1_000_000, "CSV indexer: 1 GB synthetic search_column via mmap, max_suffix_length=32, end-to-end build + query"),
    "c5": (100_000_000, 1_000_000, "max_suffix_length sweep 8/16/32/64 and query-length sweep 3-64 bytes"),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--max-suffix-length", type=int, default=L_DEFAULT)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--exchange", default="p2p", choices=["p2p", "nccl"],
                    help="global mode: fused split + peer stores through CUDA IPC (default) or split + NCCL all-to-all")
    ap.add_argument("--sharded-sa", action="store_true",
                    help="native driver: do not replicate the finished array (SA_FLAG_SHARDED_SA): every rank keeps its key "
                         "range, queries read the other ranges from the peers over NVLink")
    ap.add_argument("--replicate", default="fused", choices=["fused", "allgather"],
                    help="N > 1: how every rank gets the complete array: the finish kernel stores into all ranks' arrays "
                         "(peer stores over NVLink), or NCCL all-gather of the slices + copy")
    ap.add_argument("--chars-per-gpu", type=int, default=0, help="override the c2 text size per GPU (scale runs)")
    ap.add_argument("--queries-per-gpu", type=int, default=0, help="override the query batch size per GPU")
    ap.add_argument("--verify-single", action="store_true",
                    help="N > 1: after the timed steps rank 0 also builds the same text on ONE GPU and the arrays are "
                         "compared bit for bit; every rank's replica is compared through a digest")
    ap.add_argument("--verify-properties", action="store_true",
                    help="N > 1: size-independent checks of the global array on every rank (for n > 2^31, where no "
                         "single-GPU build exists): permutation of [0, n), first-8-byte keys non-decreasing over the "
                         "ranks, replicas identical by digest")
    ap.add_argument("--share-builder", action="store_true",
                    help="global mode: the host-buffer leg reuses the device leg's index (half the scratch; for scale runs)")
    ap.add_argument("--driver", default="native", choices=["native", "python"],
                    help="global mode: the library's own driver (sa_shard_build: one call per rank, node-local communicator, "
                         "every rank uploads only its slice of the text) or the round-1 Python loop over torch.distributed")
    ap.add_argument("--mode", default="global", choices=["global", "partition"],
                    help="N > 1: one global index built as a sample sort (default) or one independent "
                         "partition per GPU (the reference's own partitioning, suffix_array.pyx:148-180)")
    return ap.parse_args()


OVERRIDE = {}
SA_D2H = {}


def make_workload(name: str, rank: int):
    """-> (text u8, qbytes u8, qoff u64, needs_lowercase).  Seeds follow SURVEY 8(d); rank r of a
    multi-GPU run gets its own partition (seed + r)."""
    from suffixarray_b200 import workloads as W
    n_chars, nq, _ = WORKLOADS[name]
    n_chars = OVERRIDE.get("chars") or n_chars
    nq = OVERRIDE.get("queries") or nq
    if name == "c1":
        text, starts = W.company_names_joined(1_000_000, seed=1001 + rank, upper=True)
        qs = W.whole_name_queries(text, starts, nq, seed=1002 + rank)
        lens = np.fromiter((len(q) for q in qs), dtype=np.uint64, count=len(qs))
        qoff = np.zeros(len(qs) + 1, dtype=np.uint64)
        np.cumsum(lens, out=qoff[1:])
        qbytes = np.frombuffer(b"".join(qs), dtype=np.uint8).copy()
        return text, qbytes, qoff, True
    CH = 250_000_000                                   # scale runs: generated in pieces (the generator's index arrays are 8 B per char)
    if n_chars <= CH:
        text = W.english_documents(n_chars, seed=2001 + rank)
    else:
        parts, left, k = [], n_chars, 0
        while left > 0:
            take = min(CH, left) - (1 if left > CH else 0)
            parts.append(W.english_documents(take, seed=2001 + rank + 1000 * k))
            left -= take
            if left > 0:
                parts.append(np.frombuffer(b"\n", np.uint8))
                left -= 1
            k += 1
        text = np.concatenate(parts)
    qbytes, qoff = W.substring_queries(text, nq, seed=2002 + rank)
    return text, qbytes, qoff, False


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.proc = None
        self.lines: list[str] = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            parts = [p.strip() for p in ln.split(",")]
            if len(parts) < 7:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for nm, v in zip(names, parts[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def load_ncu_traffic():
    """Per-launch DRAM bytes of the onesweep kernel from the committed `ncu --set full` capture
    (profiles/ncu_traffic.json, written by profiles/summarise.py); None when absent."""
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            return json.load(f)
    except Exception:
        return None


def dist_setup(args):
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    if args.gpus > 1 and world == 1:
        raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run (one process per GPU)")
    return world, rank, local


def run_ours(args):
    import torch
    import torch.distributed as dist
    from suffixarray_b200 import _lib as S

    world, rank, local = dist_setup(args)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (libsa_cuda has no CPU fallback)")
    torch.cuda.set_device(local)
    if world > 1:
        # the "NCCL version ..." banner is written to fd 1 when the first communicator comes up: point
        # fd 1 at stderr until then, so that stdout carries the ONE JSON line and nothing else
        sys.stdout.flush()
        saved = os.dup(1)
        os.dup2(2, 1)
        try:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
            dist.barrier()
            torch.cuda.synchronize()
        finally:
            sys.stdout.flush()
            os.dup2(saved, 1)
            os.close(saved)
    S.lib()   # fail loudly when the extension is missing
    L = args.max_suffix_length
    if world > 1 and args.mode == "global":
        return (run_ours_global_native if args.driver == "native" else run_ours_global)(args, world, rank, local)

    text, qbytes, qoff, lower = make_workload(args.workload, rank)
    n, nq = int(text.size), int(qoff.size - 1)
    flags_dev = S.SA_FLAG_TIMING | (S.SA_FLAG_LOWERCASE if lower else 0)
    flags_e2e = S.SA_FLAG_LOWERCASE if lower else 0

    # pinned host staging (the e2e leg copies from / into these inside the timed region)
    h_text = torch.from_numpy(text).pin_memory()
    h_q = torch.from_numpy(qbytes).pin_memory()
    h_qoff = torch.from_numpy(qoff.view(np.int64)).pin_memory()
    h_start = torch.empty(nq, dtype=torch.int32, pin_memory=True)
    h_end = torch.empty(nq, dtype=torch.int32, pin_memory=True)
    h_hoff = torch.empty(nq + 1, dtype=torch.int64, pin_memory=True)
    d_text = torch.from_numpy(text).cuda()          # the `value` leg starts with the text in HBM

    idx_dev = S.Index(L, device=local)              # device-resident leg
    idx_e2e = S.Index(L, device=local)              # host-buffer leg
    lib = S.lib()
    for kv in (os.environ.get("SA_BENCH_DEBUG") or "").split(","):   # tuning experiments: key=value,...
        if "=" in kv:
            idx_dev.debug_set(kv.split("=")[0], int(kv.split("=")[1]))
    vp = S._vp

    h_hits = {"pos": None, "doc": None, "cap": 0}

    def ensure_hit_bufs(total):
        if total > h_hits["cap"]:
            cap = int(total * 1.1) + 1024
            h_hits["pos"] = torch.empty(cap, dtype=torch.int32, pin_memory=True)
            h_hits["doc"] = torch.empty(cap, dtype=torch.int32, pin_memory=True)
            h_hits["cap"] = cap

    def step():
        r = {}
        # ---- value leg: text resident in HBM, device-timed
        idx_dev.build_text_device(d_text.data_ptr(), n, flags_dev)
        st = idx_dev.stats()
        r["build_dev_ms"] = st["ms_total"]
        r["build_launches"] = st["kernel_launches"]
        r["onesweep_ms"] = st["ms_onesweep_sum"]
        r["onesweep_launches"] = st["radix_passes"]
        r["onesweep_elems"] = st["onesweep_elems"]
        r["phases"] = {k: round(st[k], 3) for k in ("ms_pack", "ms_hist", "ms_sort0", "ms_refine", "ms_finish",
                                                       "ms_docs", "ms_onesweep_sum")}
        r["rounds"] = st["rounds"]
        # ---- e2e leg: host text in, index ready (H2D inside)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        idx_e2e.build_text_ptr(h_text.data_ptr(), n, flags_e2e | S.SA_FLAG_TIMING)
        torch.cuda.synchronize()
        r["build_e2e_ms"] = (time.perf_counter() - t0) * 1e3
        r["build_launches"] += idx_e2e.stats()["kernel_launches"]
        # ---- queries through the C ABI with host buffers; the library's CUDA events give the
        #      device-only time of the same kernels.  Ranges (sa_query_ranges <-> get_substring_positions)
        #      and records (+ sa_query_hits <-> the hit loop of get_matching_records) are timed apart.
        t0 = time.perf_counter()
        S.check(lib.sa_query_ranges(idx_e2e._h, vp(h_q.data_ptr()), vp(h_qoff.data_ptr()), nq, K_HITS,
                                    vp(h_start.data_ptr()), vp(h_end.data_ptr()), vp(h_hoff.data_ptr())))
        torch.cuda.synchronize()
        r["range_e2e_ms"] = (time.perf_counter() - t0) * 1e3
        st = idx_e2e.stats()
        r["query_kernel_ms"] = st["ms_query"]
        r["query_launches"] = st["kernel_launches"]
        total = int(h_hoff[nq])
        ensure_hit_bufs(total)
        t0 = time.perf_counter()
        S.check(lib.sa_query_hits(idx_e2e._h, None, vp(h_hits["doc"].data_ptr()), h_hits["cap"]))
        torch.cuda.synchronize()
        r["query_e2e_ms"] = r["range_e2e_ms"] + (time.perf_counter() - t0) * 1e3
        st = idx_e2e.stats()
        r["query_dev_ms"] = st["ms_query"] + st["ms_hits"]
        r["query_launches"] = st["kernel_launches"]
        r["hits"] = total
        return r

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ensure_hit_bufs(1 << 20)
    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    barrier()
    t_wall0 = time.perf_counter()
    recs = [step() for _ in range(args.steps)]
    barrier()
    wall_ms = (time.perf_counter() - t_wall0) * 1e3
    clocks = sampler.stop() if rank == 0 else None

    # the D2H copy of the finished array is reported apart from the build (an index is
    # queryable without it); outside the timed region, pinned destination, best of 3
    h_sa = torch.empty(max(n, 1), dtype=torch.int32, pin_memory=True)
    d2h = []
    for _ in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        S.check(lib.sa_copy_sa_u32(idx_e2e._h, vp(h_sa.data_ptr()), n))
        d2h.append((time.perf_counter() - t0) * 1e3)
    SA_D2H["ms"] = min(d2h)
    del h_sa
    K = args.steps
    sums = torch.tensor([sum(r["build_dev_ms"] for r in recs), sum(r["query_dev_ms"] for r in recs),
                         sum(r["build_e2e_ms"] for r in recs), sum(r["query_e2e_ms"] for r in recs), wall_ms,
                         sum(r["range_e2e_ms"] for r in recs), sum(r["query_kernel_ms"] for r in recs)],
                        dtype=torch.float64, device="cuda")
    counts = torch.tensor([float(n), float(nq)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(sums, op=dist.ReduceOp.MAX)     # max over ranks
        dist.all_reduce(counts, op=dist.ReduceOp.SUM)   # units all ranks processed per step
    build_ms, query_ms, build_e2e_ms, query_e2e_ms, wall_ms = (sums / K).tolist()[:4] + [sums[4].item()]
    range_e2e_ms, range_dev_ms = (sums[5:7] / K).tolist()
    n_all, nq_all = counts.tolist()
    result = None
    if rank == 0:
        result = assemble(args, world, recs, n, nq, n_all, nq_all, build_ms, query_ms, build_e2e_ms, query_e2e_ms,
                          wall_ms, clocks, qbytes.size, text, qbytes, qoff, lower, range_e2e_ms, range_dev_ms)
    idx_dev.close()
    idx_e2e.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return result


def run_ours_global(args, world, rank, local):
    """N > 1: ONE index over the concatenation of the N per-rank workload texts (weak scaling of a
    single global index), built as a sample sort with an NCCL all-to-all; every rank then answers
    its own query batch against the adopted (replicated) array."""
    import torch
    import torch.distributed as dist
    from suffixarray_b200 import _lib as S
    from suffixarray_b200 import sharded as SH

    L = args.max_suffix_length
    dev = torch.device("cuda", local)
    text_r, qbytes, qoff, lower = make_workload(args.workload, rank)
    # all ranks need the whole text: all-gather the per-rank partitions (joined by '\n')
    part = np.concatenate([text_r, np.frombuffer(b"\n", np.uint8)]) if rank < world - 1 else text_r
    sizes = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([part.size], dtype=torch.int64, device=dev))
    sizes = [int(x.item()) for x in sizes]
    mx = max(sizes)
    pad = torch.zeros(mx, dtype=torch.uint8, device=dev)
    pad[: part.size] = torch.from_numpy(part).to(dev)
    allp = torch.empty(world * mx, dtype=torch.uint8, device=dev)
    dist.all_gather_into_tensor(allp, pad)
    d_text = torch.cat([allp[g * mx: g * mx + sizes[g]] for g in range(world)]).contiguous()
    del allp, pad
    h_text_t = d_text.cpu().pin_memory()
    h_text = h_text_t.numpy()
    n = int(h_text.size)
    nq = int(qoff.size - 1)
    h_q = torch.from_numpy(qbytes).pin_memory()
    h_qoff = torch.from_numpy(qoff.view(np.int64)).pin_memory()
    h_start = torch.empty(nq, dtype=torch.int32, pin_memory=True)
    h_end = torch.empty(nq, dtype=torch.int32, pin_memory=True)
    h_hoff = torch.empty(nq + 1, dtype=torch.int64, pin_memory=True)
    hits = {"doc": None, "pos": None, "cap": 0}
    lib, vp = S.lib(), S._vp
    builders = {"dev": None, "e2e": None}

    def build(kind):
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        slot = "dev" if args.share_builder else kind
        b, phases = SH.build_distributed(h_text, L, local, lowercase=lower, timing=True,
                                         device_text=d_text if kind == "dev" else None, builder=builders[slot],
                                         exchange=args.exchange, replicate=args.replicate)
        if builders[slot] is None:                                    # tuning experiments: key=value,... (from the 2nd build on)
            for kv in (os.environ.get("SA_BENCH_DEBUG") or "").split(","):
                if "=" in kv:
                    b.idx.debug_set(kv.split("=")[0], int(kv.split("=")[1]))
        builders[slot] = b
        torch.cuda.synchronize()
        return b, (time.perf_counter() - t0) * 1e3, phases

    def step():
        r = {}
        b, r["build_dev_ms"], r["phases"] = build("dev")
        st = b.idx.stats()
        r["onesweep_ms"], r["onesweep_launches"], r["onesweep_elems"] = st["ms_onesweep_sum"], st["radix_passes"], st["onesweep_elems"]
        r["build_launches"] = st["kernel_launches"]
        r["rounds"], r["m"], r["exchange_bytes"] = st["rounds"], b.m, b.exchange_bytes
        b2, r["build_e2e_ms"], _ = build("e2e")
        r["build_launches"] += b2.idx.stats()["kernel_launches"]
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        S.check(lib.sa_query_ranges(b2.idx._h, vp(h_q.data_ptr()), vp(h_qoff.data_ptr()), nq, K_HITS,
                                    vp(h_start.data_ptr()), vp(h_end.data_ptr()), vp(h_hoff.data_ptr())))
        torch.cuda.synchronize()
        r["range_e2e_ms"] = (time.perf_counter() - t0) * 1e3
        r["query_kernel_ms"] = b2.idx.stats()["ms_query"]
        total = int(h_hoff[nq])
        if total > hits["cap"]:
            hits["cap"] = int(total * 1.1) + 1024
            hits["doc"] = torch.empty(hits["cap"], dtype=torch.int32, pin_memory=True)
        t0 = time.perf_counter()
        S.check(lib.sa_query_hits(b2.idx._h, None, vp(hits["doc"].data_ptr()), hits["cap"]))
        torch.cuda.synchronize()
        r["query_e2e_ms"] = r["range_e2e_ms"] + (time.perf_counter() - t0) * 1e3
        st = b2.idx.stats()
        r["query_dev_ms"] = st["ms_query"] + st["ms_hits"]
        r["query_launches"] = st["kernel_launches"]
        r["hits"] = total
        return r

    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    dist.barrier()
    torch.cuda.synchronize()
    t_wall0 = time.perf_counter()
    recs = [step() for _ in range(args.steps)]
    dist.barrier()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t_wall0) * 1e3
    clocks = sampler.stop() if rank == 0 else None
    verify = None
    if args.verify_single:
        import hashlib
        mine = builders["dev"].idx.sa()
        digest = hashlib.blake2b(mine.tobytes(), digest_size=8).digest()
        dg = torch.tensor(list(digest), dtype=torch.int32, device=dev)
        all_dg = [torch.zeros_like(dg) for _ in range(world)]
        dist.all_gather(all_dg, dg)
        replicas_equal = all(bool(torch.equal(all_dg[0], x)) for x in all_dg)
        if rank == 0:
            for b in (builders["e2e"],):
                if b is not None:
                    b.close()
            builders["e2e"] = None
            single = S.Index(L, device=local)
            single.build_text_device(d_text.data_ptr(), n, (S.SA_FLAG_LOWERCASE if lower else 0) | S.SA_FLAG_TIMING)
            st1 = single.stats()
            same = bool(np.array_equal(single.sa(), mine))
            verify = {"single_gpu_equal": same, "replicas_equal": replicas_equal, "n": int(n),
                      "single_gpu_build_ms": round(st1["ms_total"], 2),
                      "single_gpu_mchars_per_s": round(n / st1["ms_total"] / 1e3, 1)}
            single.close()
        del mine
        dist.barrier()
    if args.verify_properties:
        verify = verify_properties(builders["dev"], d_text, n, dev, world)
    K = args.steps
    phase_names = list(recs[-1]["phases"].keys())
    sums = torch.tensor([sum(r["build_dev_ms"] for r in recs), sum(r["query_dev_ms"] for r in recs),
                         sum(r["build_e2e_ms"] for r in recs), sum(r["query_e2e_ms"] for r in recs), wall_ms,
                         sum(r["range_e2e_ms"] for r in recs), sum(r["query_kernel_ms"] for r in recs)] +
                        [sum(r["phases"][p] for r in recs) for p in phase_names], dtype=torch.float64, device=dev)
    dist.all_reduce(sums, op=dist.ReduceOp.MAX)
    ms = torch.tensor([float(recs[-1]["m"]), float(recs[-1]["exchange_bytes"])], dtype=torch.float64, device=dev)
    ms_all = [torch.zeros_like(ms) for _ in range(world)]
    dist.all_gather(ms_all, ms)
    result = None
    if rank == 0:
        build_ms, query_ms, build_e2e_ms, query_e2e_ms = (sums[:4] / K).tolist()
        range_e2e_ms, range_dev_ms = (sums[5:7] / K).tolist()
        phases = {p: round(v / K, 3) for p, v in zip(phase_names, sums[7:].tolist())}
        result = assemble(args, world, recs, n // world, nq, float(n), float(nq * world), build_ms, query_ms,
                          build_e2e_ms, query_e2e_ms, sums[4].item(), clocks, qbytes.size, text_r, qbytes, qoff, lower,
                          range_e2e_ms, range_dev_ms)
        result["e2e"]["h2d_bytes_per_step"] = int(n + qbytes.size + (nq + 1) * 8)     # per rank: the replicated text
        shard_sizes = [int(x[0].item()) for x in ms_all]
        xbytes = [int(x[1].item()) for x in ms_all]
        xms = phases.get("exchange", 0.0)
        result["config_detail"] = {}
        result["config_detail"]["sharding"] = ("one global index over the N concatenated partitions: sample sort, splitters from "
                                        "the all-reduced 65536-bin key histogram, one exchange of (key, pos) records, "
                                        "128-record halos for the bucket finish, " +
                                        ("each shard finished in chunks straight into the final array while the previous chunk is copied "
                                         "into every peer's array over NVLink (CUDA IPC, one stream per peer)"
                                         if args.replicate == "fused" else "NCCL all-gather of the slices"))
        result["config_detail"]["exchange"] = ("split kernel stores records straight into the destination GPU's sort buffers "
                                        "(CUDA IPC peer stores over NVLink)" if args.exchange == "p2p"
                                        else "split into send buffers + NCCL all_to_all_single")
        result["config_detail"]["n_chars_total"] = n
        result["timing"] = "wall clock between barrier + cudaDeviceSynchronize pairs, max over ranks (the build is host-driven across collectives)"
        if verify is not None:
            result["verify"] = verify
        result["multi_gpu"] = {"phases_ms_max_over_ranks": phases, "shard_records": shard_sizes,
                               "imbalance": round(max(shard_sizes) * world / max(sum(shard_sizes), 1), 4),
                               "exchange_bytes_sent_per_rank": xbytes,
                               "exchange_gbs_per_rank": round(max(xbytes) / (xms * 1e-3) / 1e9, 1) if xms > 0 else None,
                               "nvlink_peak_gbs_per_direction": 770.0}
    for b in builders.values():
        if b is not None:
            b.close()
    dist.barrier()
    dist.destroy_process_group()
    return result


def run_ours_global_native(args, world, rank, local):
    """N > 1: ONE index over the concatenation of the N per-rank workload texts (weak scaling of a single
    global index), built by the library's native driver: every rank hands sa_shard_build ITS text slice
    (resident in HBM for the `value` leg, in pinned host memory for the e2e leg) and ends up with the
    whole text and the complete array; it then answers its own query batch."""
    import torch
    import torch.distributed as dist
    from suffixarray_b200 import _lib as S

    L = args.max_suffix_length
    dev = torch.device("cuda", local)
    text_r, qbytes, qoff, lower = make_workload(args.workload, rank)
    part = np.concatenate([text_r, np.frombuffer(b"\n", np.uint8)]) if rank < world - 1 else text_r
    sizes = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(sizes, torch.tensor([part.size], dtype=torch.int64, device=dev))
    sizes = [int(x.item()) for x in sizes]
    n = int(sum(sizes))
    off = int(sum(sizes[:rank]))
    mlen = int(part.size)
    nq = int(qoff.size - 1)
    h_part = torch.from_numpy(part).pin_memory()              # e2e leg: this rank's slice, pinned host memory
    d_part = torch.from_numpy(part).to(dev)                   # value leg: the slice resident in HBM
    h_q = torch.from_numpy(qbytes).pin_memory()
    h_qoff = torch.from_numpy(qoff.view(np.int64)).pin_memory()
    h_start = torch.empty(nq, dtype=torch.int32, pin_memory=True)
    h_end = torch.empty(nq, dtype=torch.int32, pin_memory=True)
    h_hoff = torch.empty(nq + 1, dtype=torch.int64, pin_memory=True)
    hits = {"doc": None, "cap": 0}
    lib, vp = S.lib(), S._vp
    comm = S.Comm(f"/sa_bench_{os.environ.get('MASTER_PORT', '0')}_{os.getppid()}", rank, world)
    idx_dev = S.Index(L, device=local)
    idx_e2e = idx_dev if args.share_builder else S.Index(L, device=local)
    for kv in (os.environ.get("SA_BENCH_DEBUG") or "").split(","):
        if "=" in kv:
            idx_dev.debug_set(kv.split("=")[0], int(kv.split("=")[1]))
    base_flags = S.SA_FLAG_TIMING | (S.SA_FLAG_LOWERCASE if lower else 0) | (S.SA_FLAG_SHARDED_SA if args.sharded_sa else 0)

    def build(idx, ptr, flags):
        dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        phases = S.shard_build(idx, comm, ptr, off, mlen, n, flags)
        torch.cuda.synchronize()
        return (time.perf_counter() - t0) * 1e3, phases

    def step():
        r = {}
        r["build_dev_ms"], r["phases"] = build(idx_dev, d_part.data_ptr(), base_flags | S.SA_FLAG_DEVICE_TEXT)
        st = idx_dev.stats()
        r["onesweep_ms"], r["onesweep_launches"], r["onesweep_elems"] = st["ms_onesweep_sum"], st["radix_passes"], st["onesweep_elems"]
        r["build_launches"] = st["kernel_launches"]
        r["rounds"] = st["rounds"]
        r["m"], r["exchange_bytes"] = S.shard_info(idx_dev)
        r["build_e2e_ms"], r["phases_e2e"] = build(idx_e2e, h_part.data_ptr(), base_flags)
        r["build_launches"] += idx_e2e.stats()["kernel_launches"]
        torch.cuda.synchronize()
        if args.sharded_sa:
            dist.barrier()          # queries read the peers' shards: every rank's build must be complete
        t0 = time.perf_counter()
        S.check(lib.sa_query_ranges(idx_e2e._h, vp(h_q.data_ptr()), vp(h_qoff.data_ptr()), nq, K_HITS,
                                    vp(h_start.data_ptr()), vp(h_end.data_ptr()), vp(h_hoff.data_ptr())))
        torch.cuda.synchronize()
        r["range_e2e_ms"] = (time.perf_counter() - t0) * 1e3
        r["query_kernel_ms"] = idx_e2e.stats()["ms_query"]
        total = int(h_hoff[nq])
        if total > hits["cap"]:
            hits["cap"] = int(total * 1.1) + 1024
            hits["doc"] = torch.empty(hits["cap"], dtype=torch.int32, pin_memory=True)
        t0 = time.perf_counter()
        S.check(lib.sa_query_hits(idx_e2e._h, None, vp(hits["doc"].data_ptr()), hits["cap"]))
        torch.cuda.synchronize()
        r["query_e2e_ms"] = r["range_e2e_ms"] + (time.perf_counter() - t0) * 1e3
        st = idx_e2e.stats()
        r["query_dev_ms"] = st["ms_query"] + st["ms_hits"]
        r["query_launches"] = st["kernel_launches"]
        r["hits"] = total
        if args.sharded_sa:
            dist.barrier()          # nobody rebuilds while a peer still reads this build's shard
        return r

    for _ in range(args.warmup):
        step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    dist.barrier()
    torch.cuda.synchronize()
    t_wall0 = time.perf_counter()
    recs = [step() for _ in range(args.steps)]
    dist.barrier()
    torch.cuda.synchronize()
    wall_ms = (time.perf_counter() - t_wall0) * 1e3
    clocks = sampler.stop() if rank == 0 else None

    verify = None
    if args.verify_properties and args.sharded_sa:
        raise SystemExit("--verify-properties needs the replicated array (drop --sharded-sa)")
    if args.verify_single or args.verify_properties:
        import hashlib
        d_text = torch.from_numpy(idx_dev.text()).to(dev)       # every replica holds the whole text
        if args.verify_properties:
            class _B:
                idx = idx_dev
            verify = verify_properties(_B, d_text, n, dev, world)
        if args.verify_single:
            mine = idx_dev.sa()
            digest = hashlib.blake2b(mine.tobytes(), digest_size=8).digest()
            dg = torch.tensor(list(digest), dtype=torch.int32, device=dev)
            all_dg = [torch.zeros_like(dg) for _ in range(world)]
            dist.all_gather(all_dg, dg)
            replicas_equal = all(bool(torch.equal(all_dg[0], x)) for x in all_dg)
            if rank == 0:
                single = S.Index(L, device=local)
                single.build_text_device(d_text.data_ptr(), n, S.SA_FLAG_TIMING)      # the replicated text is already lower-cased
                st1 = single.stats()
                v = {"single_gpu_equal": bool(np.array_equal(single.sa(), mine)), "replicas_equal": replicas_equal, "n": int(n),
                     "single_gpu_build_ms": round(st1["ms_total"], 2),
                     "single_gpu_mchars_per_s": round(n / st1["ms_total"] / 1e3, 1)}
                verify = dict(verify or {}, **v)
                single.close()
            del mine
            dist.barrier()
        del d_text

    K = args.steps
    phase_names = list(recs[-1]["phases"].keys())
    sums = torch.tensor([sum(r["build_dev_ms"] for r in recs), sum(r["query_dev_ms"] for r in recs),
                         sum(r["build_e2e_ms"] for r in recs), sum(r["query_e2e_ms"] for r in recs), wall_ms,
                         sum(r["range_e2e_ms"] for r in recs), sum(r["query_kernel_ms"] for r in recs)] +
                        [sum(r["phases"][p] for r in recs) for p in phase_names] +
                        [sum(r["phases_e2e"][p] for r in recs) for p in phase_names], dtype=torch.float64, device=dev)
    dist.all_reduce(sums, op=dist.ReduceOp.MAX)
    ms = torch.tensor([float(recs[-1]["m"]), float(recs[-1]["exchange_bytes"])], dtype=torch.float64, device=dev)
    ms_all = [torch.zeros_like(ms) for _ in range(world)]
    dist.all_gather(ms_all, ms)
    ph_mine = torch.tensor([sum(r["phases"][p] for r in recs) / max(1, len(recs)) for p in phase_names], dtype=torch.float64, device=dev)
    ph_all = [torch.zeros_like(ph_mine) for _ in range(world)]
    dist.all_gather(ph_all, ph_mine)
    result = None
    if rank == 0:
        build_ms, query_ms, build_e2e_ms, query_e2e_ms = (sums[:4] / K).tolist()
        range_e2e_ms, range_dev_ms = (sums[5:7] / K).tolist()
        np_ = len(phase_names)
        phases = {p: round(v / K, 3) for p, v in zip(phase_names, sums[7:7 + np_].tolist())}
        phases_e2e = {p: round(v / K, 3) for p, v in zip(phase_names, sums[7 + np_:].tolist())}
        result = assemble(args, world, recs, n // world, nq, float(n), float(nq * world), build_ms, query_ms,
                          build_e2e_ms, query_e2e_ms, sums[4].item(), clocks, qbytes.size, text_r, qbytes, qoff, lower,
                          range_e2e_ms, range_dev_ms)
        result["e2e"]["h2d_bytes_per_step"] = int(mlen + qbytes.size + (nq + 1) * 8)     # per rank: ITS slice of the text
        shard_sizes = [int(x[0].item()) for x in ms_all]
        xbytes = [int(x[1].item()) for x in ms_all]
        xms = phases.get("exchange", 0.0)
        rep_bytes = max(shard_sizes) * 4 * (world - 1)
        result["config_detail"] = {
            "driver": "native: sa_shard_build, one call per rank; node-local communicator over POSIX shared memory + CUDA IPC",
            "sharding": "one global index over the N concatenated per-rank texts: every rank uploads its slice and pushes it to "
                        "the peers over NVLink; sample sort with splitters from the summed 65536-bin key histogram, bins weighted by the "
                        "sampled share of keys that repeat (those records go through the refinement rounds: shards are balanced "
                        "by expected work, not by record count); records "
                        "stored straight into the owner's sort buffers (peer stores); 128-record halos + device-side junction "
                        "LCPs; each shard finished in chunks into the final array while finished chunks are pushed into every "
                        "peer's array",
            "n_chars_total": n,
            "array": ("distributed (SA_FLAG_SHARDED_SA): every rank keeps its key range, queries read the others over NVLink"
                      if args.sharded_sa else "replicated on every GPU")}
        result["timing"] = "wall clock between barrier + cudaDeviceSynchronize pairs, max over ranks (the build spans several kernels, peer stores and barriers per rank)"
        if verify is not None:
            result["verify"] = verify
        result["multi_gpu"] = {"phases_ms_max_over_ranks": phases, "phases_e2e_ms_max_over_ranks": phases_e2e,
                               "phases_ms_by_rank": {p: [round(float(x[i]), 3) for x in ph_all] for i, p in enumerate(phase_names)},
                               "shard_records": shard_sizes,
                               "imbalance": round(max(shard_sizes) * world / max(sum(shard_sizes), 1), 4),
                               "exchange_bytes_sent_per_rank": xbytes,
                               "exchange_gbs_per_rank": round(max(xbytes) / (xms * 1e-3) / 1e9, 1) if xms > 0 else None,
                               "text_allgather": ("asynchronous: every rank's slice goes to the peers on the copy engines beside "
                                                  "prepare / pack / split / exchange; the `text` phase is the upload of the own slice, "
                                                  "the wait for the peers' slices is inside `exchange`"),
                               "text_bytes_sent_per_rank": int(mlen * (world - 1)),
                               "replicate_bytes_sent_per_rank": int(rep_bytes),
                               "finish_replicate_gbs_per_rank": (round(rep_bytes / (phases["finish+replicate"] * 1e-3) / 1e9, 1)
                                                                 if phases.get("finish+replicate", 0) > 0 else None),
                               "nvlink_peak_gbs_per_direction": 900.0}
    if idx_e2e is not idx_dev:
        idx_e2e.close()
    idx_dev.close()
    comm.barrier()
    comm.close()
    dist.barrier()
    dist.destroy_process_group()
    return result


class _DeviceU32:
    """zero-copy torch view of a raw device array (CUDA array interface)"""

    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<i4", "data": (int(ptr), False), "version": 2}


def array_properties(sa, d_text, n, chunk=1 << 25):
    """sa: int32 tensor (u32 bit patterns) of n entries, d_text: uint8 tensor of the n text bytes, both on
    one device.  -> (is a permutation of [0, n), masked first-8-byte keys non-decreasing or None for
    non-ASCII text, 2-word digest).  Chunked: scratch is 8 index words per entry of a chunk."""
    import torch
    dev = d_text.device
    seen = torch.zeros(n, dtype=torch.uint8, device=dev)
    text = torch.cat([d_text, torch.zeros(8, dtype=torch.uint8, device=dev)])
    ar8 = torch.arange(8, device=dev, dtype=torch.int64)
    shifts = 8 * (7 - ar8)
    sorted_ok, prev_last = True, None
    digest = torch.zeros(2, dtype=torch.int64, device=dev)
    ascii_only = bool((d_text < 128).all())
    for a in range(0, n, chunk):
        z = min(n, a + chunk)
        pos = sa[a:z].to(torch.int64) & 0xFFFFFFFF
        if bool((pos >= n).any()):
            return False, False if ascii_only else None, digest
        seen[pos] = 1
        weight = torch.arange(a, z, device=dev, dtype=torch.int64) % 1000003 + 1
        digest[0] += (pos * weight).sum()
        digest[1] += (pos * pos).sum()
        if ascii_only:
            b = text[pos[:, None] + ar8[None, :]].to(torch.int64)
            term = ((b == 10) | (b == 255)).to(torch.int32)
            after = (torch.cumsum(term, dim=1) - term) > 0
            b = torch.where(after, torch.zeros_like(b), b)
            key = (b << shifts[None, :]).sum(dim=1)
            sorted_ok = sorted_ok and bool((key[1:] >= key[:-1]).all())
            if prev_last is not None:
                sorted_ok = sorted_ok and bool(key[0] >= prev_last)
            prev_last = key[-1].clone()
            del b, term, after, key
        del pos, weight
    return bool(seen.all()), (sorted_ok if ascii_only else None), digest


def verify_properties(builder, d_text, n, dev, world):
    """Checks that need no reference array, on the device, in chunks:
    the array is a permutation of [0, n); K(SA[i]) <= K(SA[i+1]) for the masked big-endian key K of
    the first 8 bytes (ORDER-E refines this order, it never contradicts it); all replicas agree."""
    import torch
    import torch.distributed as dist
    from suffixarray_b200 import _lib as S
    sa = torch.as_tensor(_DeviceU32(S.lib().sa_device_sa(builder.idx._h), n), device=dev)
    perm_ok, sorted_ok, digest = array_properties(sa, d_text, n)
    all_dg = [torch.zeros_like(digest) for _ in range(world)]
    dist.all_gather(all_dg, digest)
    flags = torch.tensor([int(perm_ok), int(sorted_ok is not False)], dtype=torch.int32, device=dev)
    dist.all_reduce(flags, op=dist.ReduceOp.MIN)
    return {"n": int(n), "permutation_on_every_rank": bool(flags[0].item()),
            "first_8_byte_keys_non_decreasing_on_every_rank": bool(flags[1].item()) if sorted_ok is not None else None,
            "replicas_equal": all(bool(torch.equal(all_dg[0], x)) for x in all_dg),
            "positions_above_2^31": int(max(0, n - (1 << 31)))}


# ------------------------------------------------------------------------------ CPU reference legs
def cpu_reference_times(text, qbytes, qoff, lower, L, build_reps=1, query_sample=0, threads=0,
                        record_sample=20_000, keep=False):
    """Times the reference's own CPU path (engine.c through oracle/_ref/libsa_ref.so when present,
    else the oracle port) on the host cores.  query_sample = 0: the whole batch.  The OpenMP
    thread count is set explicitly to the cores this process may run on (torch.distributed.run
    exports OMP_NUM_THREADS=1 to its workers).  keep=True also returns what the parity check
    needs (the text as indexed, the reference's array and ranges).  -> dict."""
    from oracle import pyoracle as O
    nq_all = int(qoff.size - 1)
    nqs = min(nq_all, query_sample) if query_sample else nq_all
    low = O.oracle_lowercase(text) if lower else text
    n = int(low.size)
    kind = "reference" if O.have_ref() else "port"
    cores = (threads or O.ref_use_all_cores()) if kind == "reference" else 1
    rec = {"record_sample": 0, "record_s": 0.0, "records": 0}
    kept = None
    if kind == "reference":
        O.ref_lib().ref_set_threads(cores)
        # The CPU arm indexes '\n' + text (one empty leading record): get_matching_records scans back
        # to the previous '\n' with a uint32 offset (engine.c:1189-1191) and leaves the buffer for any
        # hit in record 0.  [ text' | 64 zero bytes | '\n' | slack ] keeps the forward scan
        # (:1194-1205) and the after-the-record memcpy (:1208) inside the buffer too.
        GUARD = 1 << 16
        n = n + 1
        buf = np.zeros(n + 64 + 1 + GUARD, dtype=np.uint8)
        buf[0] = 10
        buf[1:n] = low
        buf[n + 64] = 10
        tp = buf
        sa = np.empty(n, dtype=np.uint32)
        best = float("inf")
        for _ in range(build_reps):
            t0 = time.perf_counter()
            O.ref_build_padded(tp, n, L, sa)
            best = min(best, time.perf_counter() - t0)
        # query batch: NUL-terminated C strings, an omp parallel-for over the batch
        lens = (qoff[1:nqs + 1] - qoff[:nqs]).astype(np.int64)
        zoff = np.zeros(nqs + 1, dtype=np.uint64)
        np.cumsum(lens + 1, out=zoff[1:].view(np.int64))
        zbytes = np.zeros(int(zoff[nqs]), dtype=np.uint8)
        src = qbytes[: int(qoff[nqs])]
        dst_idx = np.arange(src.size, dtype=np.int64) + np.repeat(np.arange(nqs, dtype=np.int64), lens)
        zbytes[dst_idx] = src
        start = np.full(nqs, 0xFFFFFFFF, dtype=np.uint32)
        end = np.full(nqs, 0xFFFFFFFF, dtype=np.uint32)
        skipped = np.zeros(nqs, dtype=np.uint8)
        import ctypes as C
        rl = O.ref_lib()
        u32p, u64p, u8p = C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_uint8)
        t0 = time.perf_counter()
        rl.ref_query_batch(tp.ctypes.data, sa.ctypes.data_as(u32p), n, L, zbytes.ctypes.data,
                           zoff.ctypes.data_as(u64p), nqs, start.ctypes.data_as(u32p), end.ctypes.data_as(u32p),
                           skipped.ctypes.data_as(u8p), cores)
        tq = time.perf_counter() - t0
        # record path: get_matching_records (bisection + newline scans + malloc/memcpy per hit), smaller sample
        nrec = min(nqs, record_sample)
        hits = np.zeros(max(nrec, 1), dtype=np.uint32)
        t0 = time.perf_counter()
        total = rl.ref_query_records_batch(tp.ctypes.data, sa.ctypes.data_as(u32p), n, L, zbytes.ctypes.data,
                                           zoff.ctypes.data_as(u64p), nrec, K_HITS, hits.ctypes.data_as(u32p), cores)
        rec = {"record_sample": nrec, "record_s": time.perf_counter() - t0, "records": int(total)}
        if keep:
            kept = {"text": tp[:n], "sa": sa, "start": start, "end": end, "skipped": skipped, "nq": nqs}
    else:
        t0 = time.perf_counter()
        sa = O.oracle_build(low, L)
        best = time.perf_counter() - t0
        from suffixarray_b200 import workloads as W
        qs = W.unpack_queries(qbytes[: int(qoff[nqs])], qoff[: nqs + 1])
        t0 = time.perf_counter()
        ws, we = O.oracle_query(low, sa, L, qs)
        tq = time.perf_counter() - t0
        if keep:
            kept = {"text": low, "sa": sa, "start": ws, "end": we, "skipped": np.zeros(nqs, np.uint8), "nq": nqs}
    out = {"kind": kind, "cores": int(cores), "n": n, "build_s": best, "build_mchars_s": n / best / 1e6,
           "query_sample": nqs, "query_s": tq, "queries_per_s": nqs / tq}
    out.update(rec)
    out["record_queries_per_s"] = rec["record_sample"] / rec["record_s"] if rec["record_s"] > 0 else None
    if keep:
        out["kept"] = kept
    return out


def parity_check(kept, qbytes, qoff, L, device):
    """The bench's own parity gate: the SAME text the CPU arm indexed is built on the GPU (outside the
    timed region) and the arrays are compared entry by entry; so are the rank ranges of every query
    the reference answered without its out-of-bounds wrap, and the hit lists of a 5 000-query sample."""
    from oracle import pyoracle as O
    from suffixarray_b200 import _lib as S
    idx = S.Index(L, device=device)
    idx.build_text(kept["text"])
    sa_equal = bool(np.array_equal(idx.sa(), kept["sa"]))
    nq = kept["nq"]
    s, e, off = idx.query_ranges_packed(qbytes[: max(int(qoff[nq]), 1)], qoff[: nq + 1], k=K_HITS)
    keep = kept["skipped"] == 0
    ranges_equal = bool(np.array_equal(s[keep], kept["start"][keep]) and np.array_equal(e[keep], kept["end"][keep]))
    pos, doc = idx.query_hits(int(off[-1]))
    pick = np.flatnonzero(keep)[:5000]
    woff, wpos, wdoc = O.oracle_hits(kept["text"], kept["sa"], kept["start"][pick], kept["end"][pick], K_HITS)
    hits_equal = True
    for t, q in enumerate(pick.tolist()):
        a, b, wa, wb = int(off[q]), int(off[q + 1]), int(woff[t]), int(woff[t + 1])
        if b - a != wb - wa or not np.array_equal(pos[a:b], wpos[wa:wb]) or not np.array_equal(doc[a:b], wdoc[wa:wb]):
            hits_equal = False
            break
    idx.close()
    return {"sa_equal": sa_equal, "n": int(kept["sa"].size), "ranges_equal": ranges_equal,
            "queries_compared": int(keep.sum()), "queries_skipped_reference_wrap": int((~keep).sum()),
            "hit_lists_equal": hits_equal, "hit_lists_compared": int(pick.size),
            "against": "oracle/_ref (engine.c)" if O.have_ref() else "oracle port"}


def make_config(args, world, n, nq):
    """The workload description, identical in both arms (`ours` and `--impl reference`)."""
    glob = world > 1 and args.mode == "global"
    return {"workload": WORKLOADS[args.workload][2], "name": args.workload, "n_chars_per_gpu": int(n),
            "queries_per_gpu": int(nq), "max_suffix_length": args.max_suffix_length, "k": K_HITS,
            "sharding": ("one global index over the N concatenated per-GPU texts (sample sort)" if glob else
                         "one independent partition per GPU (suffix_array.pyx:148-180)"),
            "l2": "inputs larger than L2 (keys 8n + positions 4n = 1.2 GB per pass vs 126 MB L2)"}


def assemble(args, world, recs, n, nq, n_all, nq_all, build_ms, query_ms, build_e2e_ms, query_e2e_ms, wall_ms,
             clocks, qbytes_len, text, qbytes, qoff, lower, range_e2e_ms, range_dev_ms):
    K = args.steps
    peak, peak_src = load_peaks()
    os_ms = sum(r["onesweep_ms"] for r in recs)
    os_launches = sum(r["onesweep_launches"] for r in recs)
    os_elems = sum(r["onesweep_elems"] for r in recs)
    os_bytes_per_launch = 24.0 * os_elems / max(os_launches, 1)
    os_avg_ms = os_ms / max(os_launches, 1)
    achieved = os_bytes_per_launch / (os_avg_ms * 1e-3) / 1e9 if os_avg_ms > 0 else 0.0
    traffic = load_ncu_traffic()
    hits = recs[-1]["hits"]
    h2d = n + qbytes_len + (nq + 1) * 8
    d2h = nq * 8 + (nq + 1) * 8 + hits * 4
    out = {
        "metric": METRIC,
        "value": round(n_all / (build_ms * 1e-3) / 1e6, 2),
        "unit": "Mchars/s (SA build)",
        "queries_per_s": round(nq_all / (range_dev_ms * 1e-3), 1),
        "record_queries_per_s": round(nq_all / (query_ms * 1e-3), 1),
        "queries_note": "queries_per_s = rank ranges (sa_query_ranges <-> get_substring_positions, engine.c:869); "
                        "record_queries_per_s adds the first min(k, count) hits per query as record ids "
                        "(sa_query_hits <-> the hit loop of get_matching_records, engine.c:1185-1211)",
        "n_gpus": world, "steps": K, "warmup": args.warmup,
        "ms_per_step": round(build_ms + query_ms, 3),
        "build_ms_per_step": round(build_ms, 3), "query_ms_per_step": round(query_ms, 3),
        "wall_ms_timed_region": round(wall_ms, 1),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8 text / u64 keys / u32 positions (integer compare + scatter)",
        "data": "synthetic",
        "config": make_config(args, world, n, nq),
        "hits_per_batch": hits,
        "e2e": {"value": round(n_all / (build_e2e_ms * 1e-3) / 1e6, 2), "unit": "Mchars/s (SA build)",
                "queries_per_s": round(nq_all / (range_e2e_ms * 1e-3), 1),
                "record_queries_per_s": round(nq_all / (query_e2e_ms * 1e-3), 1),
                "build_ms": round(build_e2e_ms, 3), "range_query_ms": round(range_e2e_ms, 3),
                "record_query_ms": round(query_e2e_ms, 3),
                "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "api": "sa_build_text + sa_query_ranges + sa_query_hits (pinned host buffers)"},
        "gpu_launches": int(sum(r["build_launches"] + r["query_launches"] for r in recs)),
        "roofline": {"kernel": "onesweep4_kernel (radix scatter pass)", "bound": "hbm",
                     "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(achieved / peak, 4), "peak_source": peak_src,
                     "bytes_per_launch": int(os_bytes_per_launch), "avg_launch_ms": round(os_avg_ms, 4),
                     "launches_per_step": os_launches / K,
                     "traffic": (round((traffic or {}).get("onesweep_dram_bytes_per_record") * os_elems / max(os_launches, 1))
                                 if (traffic or {}).get("onesweep_dram_bytes_per_record") else None),
                     "traffic_source": "profiles/ncu_traffic.json: dram__bytes_read+write per record of the captured "
                                       "100 M-record launches x records per launch here"},
        "roofline_build_model": {"bytes": 213 * n, "achieved": round(213.0 * n / (build_ms * 1e-3) / 1e9, 1),
                                 "frac": round(213.0 * n / (build_ms * 1e-3) / 1e9 / peak, 4),
                                 "note": "SURVEY 8(d) B_build = 213 n bytes over the whole device build"},
        "phases_ms_last_step": recs[-1]["phases"],
        "refine_rounds": recs[-1]["rounds"],
        "query_kernel_ms": round(range_dev_ms, 3),
        "clocks": clocks,
    }
    if SA_D2H.get("ms"):
        out["e2e"]["sa_d2h_ms"] = round(SA_D2H["ms"], 3)
        out["e2e"]["sa_d2h_bytes"] = int(4 * n)
        out["e2e"]["build_plus_sa_d2h_mchars_per_s"] = round(n_all / ((build_e2e_ms + SA_D2H["ms"]) * 1e-3) / 1e6, 2)
        out["e2e"]["sa_d2h_note"] = ("sa_copy_sa_u32 into a pinned buffer, per GPU, outside the timed region (SURVEY 8(d): "
                                     "text H2D is inside `build_ms`, the array's D2H is reported apart and combined here)")
    # query kernel against its roofline (SURVEY 8(d) B_query): per query 2*ceil(log2 n) probes, each an
    # SA entry (4 B) + the compared window (c = min(|q|, L) bytes), + the 16-byte result; the sector
    # figure counts what the memory system moves for them (32 B per SA entry, 32*(ceil(c/32)+1) per window)
    if qoff is not None and range_dev_ms > 0:
        import math
        n_index = n * world if (world > 1 and args.mode == "global") else n
        probes = 2 * math.ceil(math.log2(max(n_index, 2)))
        lens = np.minimum((qoff[1:] - qoff[:-1]).astype(np.int64), args.max_suffix_length)
        useful = float(probes) * float((4 + lens).sum()) + 16.0 * lens.size
        sector = float(probes) * float((32 + 32 * ((lens + 31) // 32 + 1)).sum())
        t = range_dev_ms * 1e-3
        out["roofline_query"] = {"kernel": "query_thread_kernel (two bisections per query, one query per thread)",
                                 "bound": "hbm", "unit": "GB/s", "peak": peak,
                                 "achieved": round(useful / t / 1e9, 1), "frac": round(useful / t / 1e9 / peak, 4),
                                 "sector_achieved": round(sector / t / 1e9, 1),
                                 "sector_frac": round(sector / t / 1e9 / peak, 4),
                                 "probes_per_query": probes, "bytes_per_query": round(useful / lens.size, 1),
                                 "sector_bytes_per_query": round(sector / lens.size, 1), "launch_ms": round(range_dev_ms, 4),
                                 "note": "achieved = useful bytes (probes*(4+c)+16) / kernel time; sector_* = 32-byte "
                                         "sectors those dependent random accesses touch"}
    if not args.no_cpu_baseline and text is not None:
        cb = cpu_reference_times(text, qbytes, qoff, lower, args.max_suffix_length, keep=True)
        out["cpu_baseline"] = {"value": round(cb["build_mchars_s"], 2), "unit": "Mchars/s (SA build)",
                               "queries_per_s": round(cb["queries_per_s"], 1),
                               "record_queries_per_s": round(cb["record_queries_per_s"], 1) if cb["record_queries_per_s"] else None,
                               "cores": cb["cores"], "kind": cb["kind"],
                               "sample": ("rank 0's text and batch; " if world > 1 else "") +
                                         f"build: the full {cb['n']}-char text once ({cb['build_s']:.2f} s); "
                                         f"queries: all {cb['query_sample']} of the batch ({cb['query_s']:.2f} s), "
                                         "get_substring_positions under an omp parallel-for; records: first "
                                         f"{cb['record_sample']} queries through get_matching_records, k = {K_HITS} "
                                         f"({cb['records']} records, {cb['record_s']:.2f} s)"}
        try:
            pc = parity_check(cb.pop("kept"), qbytes, qoff, args.max_suffix_length, int(os.environ.get("LOCAL_RANK", "0")))
            out["parity_checked"] = bool(pc["sa_equal"] and pc["ranges_equal"] and pc["hit_lists_equal"])
            out["parity"] = pc
        except Exception as ex:      # the bench line must still be printed
            out["parity_checked"] = False
            out["parity"] = {"error": repr(ex)[:300]}
        if cb["kind"] == "reference" and world == 1:
            # BASELINE.json's other CPU arm: the vendored libsais (a FULL suffix array, not ORDER-E: timing only)
            from oracle import pyoracle as O
            low = (O.oracle_lowercase(text) if lower else text)[:50_000_000]      # bounded sample
            t0 = time.perf_counter()
            O.ref_libsais(low, threads=cb["cores"])
            tl = time.perf_counter() - t0
            out["cpu_baseline"]["libsais_omp"] = {"value": round(low.size / tl / 1e6, 2), "unit": "Mchars/s (full SA, untruncated)",
                                                  "cores": cb["cores"], "seconds": round(tl, 2), "sample_chars": int(low.size),
                                                  "note": "libsais_omp from the reference tree (SURVEY 8(d) baseline B); not a parity oracle"}
    return out


def run_reference(args):
    """The reference's own CPU implementation (engine.c) on the host cores: rank 0 only, all the
    cores this process may run on.  Each step = the per-GPU workload of the GPU arm's config: the
    whole text built once + the whole query batch (+ a 20 000-query sample of the record path).  At
    N > 1 that is rank 0's share of the N-GPU job (a 1/N sample of the global text)."""
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if rank != 0:
        return None
    text, qbytes, qoff, lower = make_workload(args.workload, 0)
    n, nq = int(text.size), int(qoff.size - 1)
    recs = []
    for i in range(args.warmup + args.steps):
        cb = cpu_reference_times(text, qbytes, qoff, lower, args.max_suffix_length)
        if i >= args.warmup:
            recs.append(cb)
    K = len(recs)
    bs = sum(r["build_s"] for r in recs) / K
    qs = sum(r["query_s"] for r in recs) / K
    v = recs[0]["n"] / bs / 1e6
    rqs = [r["record_queries_per_s"] for r in recs if r.get("record_queries_per_s")]
    rq = round(sum(rqs) / len(rqs), 1) if rqs else None
    sample = (f"per step: construct_truncated_suffix_array over the whole {recs[0]['n']}-char text + "
              f"get_substring_positions for all {recs[0]['query_sample']} queries of the batch (omp parallel-for) + "
              f"get_matching_records for the first {recs[0]['record_sample']} queries" +
              (f"; N = {world}: rank 0's share of the {world}-GPU job" if world > 1 else ""))
    return {"impl": "reference", "metric": METRIC, "value": round(v, 2), "unit": "Mchars/s (SA build)",
            "queries_per_s": round(recs[0]["query_sample"] / qs, 1), "record_queries_per_s": rq, "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": round((bs + qs) * 1e3, 2), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8 text / u64 keys / u32 positions (integer compare + scatter)", "data": "synthetic",
            "config": make_config(args, world, n, nq),
            "cpu_baseline": {"value": round(v, 2), "unit": "Mchars/s (SA build)", "kind": recs[0]["kind"],
                             "cores": recs[0]["cores"], "queries_per_s": round(recs[0]["query_sample"] / qs, 1),
                             "record_queries_per_s": rq, "sample": sample},
            "e2e": {"value": round(v, 2), "unit": "Mchars/s (SA build)", "queries_per_s": round(recs[0]["query_sample"] / qs, 1),
                    "record_queries_per_s": rq, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}


def run_sweep_workload(args):
    """--workload c3 | c5: BASELINE configs 3 and 5 on one GPU through scripts/sweeps.py, every GPU number
    with the reference's CPU number beside it (engine.c on all host cores; the CSV path end to end through
    construct_..._mmap_full and get_matching_records_file).  One JSON line in the bench's shape."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("sweeps", os.path.join(ROOT, "scripts", "sweeps.py"))
    sw = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(sw)

    class A:
        chars = OVERRIDE.get("chars") or 100_000_000
        queries = OVERRIDE.get("queries") or 1_000_000
        csv_mb = int(os.environ.get("SA_BENCH_CSV_MB", "1024"))
        no_cpu = args.no_cpu_baseline
    out = {}
    if args.workload == "c3":
        sw.section_c3(A, out)
        c = out["c3"]
        line = {"metric": METRIC, "value": c["mchars_per_s_e2e"], "unit": "Mchars/s (CSV file -> queryable index, end to end)",
                "n_gpus": 1, "higher_is_better": True, "data": "synthetic", "dtype": "u8 / u64 keys / u32 positions",
                "config": {"workload": WORKLOADS["c3"][2], "name": "c3"},
                "cpu_baseline": dict(c.get("cpu_engine_c", {}), kind="reference", value=c.get("cpu_engine_c", {}).get("mchars_per_s_e2e")),
                "sweep": out}
    else:
        sw.section_c5(A, out)
        r32 = [r for r in out["c5_L_sweep"]["rows"] if r["max_suffix_length"] == 32][0]
        line = {"metric": METRIC, "value": r32["mchars_per_s"], "unit": "Mchars/s (SA build, L = 32 row of the sweep)", "n_gpus": 1,
                "higher_is_better": True, "data": "synthetic", "dtype": "u8 / u64 keys / u32 positions",
                "config": {"workload": WORKLOADS["c5"][2], "name": "c5"},
                "cpu_baseline": dict(r32["cpu_engine_c"], kind="reference", value=r32["cpu_engine_c"]["mchars_per_s"]),
                "sweep": out}
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    if args.chars_per_gpu:
        OVERRIDE["chars"] = args.chars_per_gpu
    if args.queries_per_gpu:
        OVERRIDE["queries"] = args.queries_per_gpu
    if args.workload in ("c3", "c5"):
        return run_sweep_workload(args)
    res = run_reference(args) if args.impl == "reference" else run_ours(args)
    if res is not None:
        if OVERRIDE and "config" in res:
            res["config"]["workload"] += f" -- SCALED RUN: {OVERRIDE} per GPU instead of the named sizes"
        print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
