#!/usr/bin/env python
"""BASELINE.json configs beside the headline one, on 1 GPU (run on the GPU box):

  c1  1M company names (~20 MB), L=32: build + 10k whole-name query_records (device + host-buffer legs)
  c3  CSV indexer end to end: synthetic `name,city,score` file, search column `name`, through
      sa_build_csv (mmap scan -> pinned buffers -> async H2D -> build -> remap) + a query batch
  c5  max_suffix_length sweep 8/16/32/64 on the C2 text, and query-length sweep 3..64 bytes

Each section is checked against a size-independent property (sortedness sample, hit verification)
so a fast wrong answer cannot slip in; parity proper is tests/.  Writes one JSON document.

    python scripts/sweeps.py --out gpurun_out/sweeps.json [--csv-mb 1024] [--chars 100000000]
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def sync():
    import torch
    torch.cuda.synchronize()


def timed(fn, reps=3, warm=1):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(reps):
        sync()
        t0 = time.perf_counter()
        fn()
        sync()
        ts.append((time.perf_counter() - t0) * 1e3)
    return float(np.median(ts))


def check_hits_sample(text: np.ndarray, sa: np.ndarray, qbytes, qoff, start, end, L: int, sample=2000, seed=7):
    """every sampled rank inside [start, end) must hold a suffix that starts with the query
    (truncated to L), and the ranks just outside must not."""
    rng = np.random.default_rng(seed)
    nq = qoff.size - 1
    tb = text.tobytes()
    qb = qbytes.tobytes()
    bad = 0
    for i in rng.integers(0, nq, size=min(sample, nq)):
        q = qb[int(qoff[i]):int(qoff[i + 1])][:L]
        s, e = int(start[i]), int(end[i])
        for r in {s, (s + e) // 2, e - 1} if e > s else ():
            p = int(sa[r])
            if tb[p:p + len(q)] != q:
                bad += 1
    return bad


def section_c5(args, out):
    import torch
    from suffixarray_b200 import _lib as S
    from suffixarray_b200 import workloads as W
    text = W.english_documents(args.chars, seed=2001)
    n = text.size
    d_text = torch.from_numpy(text).cuda()
    rows = []
    for L in (8, 16, 32, 64):
        idx = S.Index(L, device=0)
        ms = timed(lambda: idx.build_text_device(d_text.data_ptr(), n, S.SA_FLAG_TIMING), reps=3, warm=2)
        st = idx.stats()
        rows.append({"max_suffix_length": L, "build_ms": round(st["ms_total"], 3), "wall_ms": round(ms, 3),
                     "mchars_per_s": round(n / st["ms_total"] / 1e3, 1), "refine_rounds": st["rounds"],
                     "radix_passes": st["radix_passes"], "finish_ms": round(st["ms_finish"], 3)})
        if L != 32:
            idx.close()
        else:
            idx32 = idx
    out["c5_L_sweep"] = {"workload": f"{n} chars synthetic English (C2 text)", "rows": rows}
    # query-length sweep against the L=32 index, device-timed kernel (library CUDA events) + host-buffer wall time
    sa = idx32.sa()
    qrows = []
    nq = args.queries
    for m in (3, 4, 6, 8, 12, 16, 24, 32, 48, 64):
        qbytes, qoff = W.fixed_length_queries(text, nq, m, seed=5000 + m)
        res = {}

        def run():
            res["r"] = idx32.query_ranges_packed(qbytes, qoff, k=1000, want_offsets=True)
        ms = timed(run, reps=3, warm=1)
        st = idx32.stats()
        start, end, hoff = res["r"]
        bad = check_hits_sample(text, sa, qbytes, qoff, start, end, 32, sample=300)
        qrows.append({"query_bytes": m, "kernel_ms": round(st["ms_query"], 4),
                      "queries_per_s_device": round(nq / st["ms_query"] * 1e3, 0),
                      "queries_per_s_host_buffers": round(nq / ms * 1e3, 0), "total_hits_k1000": int(hoff[-1]),
                      "sample_mismatches": bad})
    out["c5_query_length_sweep"] = {"queries": nq, "index": "L=32, C2 text", "rows": qrows}
    idx32.close()


def section_c1(args, out):
    import torch
    from suffixarray_b200 import _lib as S
    from suffixarray_b200 import workloads as W
    text, starts = W.company_names_joined(1_000_000, seed=1001, upper=True)
    qs = W.whole_name_queries(text, starts, 10_000, seed=1002)
    n = text.size
    idx = S.Index(32, device=0)
    ms = timed(lambda: idx.build_text(text, lowercase=True, timing=True), reps=3, warm=1)
    st = idx.stats()
    res = {}

    def q():
        s, e, off = idx.query_ranges(qs, k=1000)
        res["hits"] = idx.query_hits(int(off[-1]))
        res["off"] = off
    qms = timed(q, reps=3, warm=1)
    out["c1"] = {"workload": f"1M company names, {n} chars, L=32, 10k whole-name queries, k=1000",
                 "build_device_ms": round(st["ms_total"], 3), "build_host_buffers_ms": round(ms, 3),
                 "mchars_per_s_e2e": round(n / ms / 1e3, 1), "code_bits": st["code_bits"], "refine_rounds": st["rounds"],
                 "query_records_ms_host_buffers": round(qms, 3), "record_queries_per_s_e2e": round(10_000 / qms * 1e3, 0),
                 "hits": int(res["off"][-1])}
    idx.close()


def write_csv(path: str, target_mb: int, seed: int = 3001):
    """`name,city,score` rows until the name column holds about target_mb MB.  Assembled with numpy
    (a Python row loop would cost minutes of box time at 40 M rows); quoted cells are left to the
    parity tests (tests/test_gpu_dropin.py, tests/golden)."""
    from suffixarray_b200 import workloads as W
    suffix = np.frombuffer(b",riverton,0421", dtype=np.uint8)
    ns = suffix.size
    col_bytes = 0
    total_rows = 0
    chunk = 0
    with open(path, "wb") as f:
        f.write(b"name,city,score\n")
        while col_bytes < target_mb * 1_000_000:
            rows = 2_000_000
            text, _ = W.company_names_joined(rows, seed=seed + 17 * chunk, upper=False)
            text = np.concatenate([text, np.frombuffer(b"\n", np.uint8)])
            nl = text == 10
            k = np.cumsum(nl, dtype=np.int64)
            dst = np.arange(text.size, dtype=np.int64) + ns * k
            out = np.empty(text.size + ns * int(k[-1]), dtype=np.uint8)
            out[dst] = text
            at = dst[nl] - ns
            out[(at[:, None] + np.arange(ns)[None, :]).ravel()] = np.tile(suffix, at.size)
            out.tofile(f)
            col_bytes += text.size
            total_rows += rows
            chunk += 1
    return total_rows, os.path.getsize(path)


def section_c3(args, out):
    from suffixarray_b200 import _lib as S
    from suffixarray_b200 import workloads as W
    d = tempfile.mkdtemp(prefix="sa_c3_")
    path = os.path.join(d, "c3.csv")
    t0 = time.perf_counter()
    rows, fbytes = write_csv(path, args.csv_mb)
    gen_s = time.perf_counter() - t0
    idx = S.Index(32, device=0)
    walls = []
    for rep in range(3):
        sync()
        t0 = time.perf_counter()
        idx.build_csv(path, "name", timing=True)
        sync()
        walls.append((time.perf_counter() - t0) * 1e3)
    st = idx.stats()
    n = idx.n
    text = idx.text()
    sa_text_order = None
    # queries: substrings of the indexed column text
    qbytes, qoff = W.substring_queries(text, args.queries, seed=3002)
    res = {}

    def q():
        res["r"] = idx.query_ranges_packed(qbytes, qoff, k=1000, want_offsets=True)
    qms = timed(q, reps=3, warm=1)
    stq = idx.stats()
    out["c3"] = {"workload": f"CSV {fbytes / 1e6:.0f} MB, {rows} rows, search column `name` = {n} chars, L=32",
                 "generate_s": round(gen_s, 1),
                 "build_e2e_ms_runs": [round(w, 1) for w in walls], "build_e2e_ms": round(min(walls[1:]), 1),
                 "mchars_per_s_e2e": round(n / min(walls[1:]) / 1e3, 1),
                 "file_mb_per_s_e2e": round(fbytes / min(walls[1:]) / 1e3, 1),
                 "device_build_ms": round(st["ms_total"], 2), "csv_scan_ms": round(st.get("ms_csv_scan", 0.0), 2),
                 "phases_ms": {k: round(v, 2) for k, v in st.items() if k.startswith("ms_")},
                 "range_queries_per_s_device": round(args.queries / stq["ms_query"] * 1e3, 0),
                 "range_queries_per_s_host_buffers": round(args.queries / qms * 1e3, 0)}
    # record queries: hits come back as (file offset, row of the file) from the device row table
    total = int(res["r"][2][-1])

    def qh():
        res["h"] = idx.query_hits(total)
    hms = timed(qh, reps=3, warm=1)
    sth = idx.stats()
    pos, row = res["h"]
    out["c3"].update({"hits_per_batch": total, "hits_ms_device": round(sth["ms_hits"], 3), "hits_ms_host_buffers": round(hms, 3),
                      "record_queries_per_s_device": round(args.queries / (stq["ms_query"] + sth["ms_hits"]) * 1e3, 0),
                      "rows_in_range": bool(total == 0 or int(row.max()) <= rows)})
    idx.close()
    # three search columns from one read of the file (sa_build_csv_columns)
    cols = ["name", "city", "score"]
    idxs = [S.Index(32, device=0) for _ in cols]
    walls3 = []
    for rep in range(2):
        sync()
        t0 = time.perf_counter()
        S.Index.build_csv_columns(idxs, path, cols, timing=True)
        sync()
        walls3.append((time.perf_counter() - t0) * 1e3)
    out["c3"]["three_columns_one_call"] = {
        "columns": cols, "chars": [ix.n for ix in idxs], "e2e_ms_runs": [round(w, 1) for w in walls3],
        "device_build_ms": [round(ix.stats()["ms_total"], 2) for ix in idxs],
        "csv_scan_ms": [round(ix.stats()["ms_csv_scan"], 2) for ix in idxs],
        "h2d_ms": [round(ix.stats()["ms_h2d"], 2) for ix in idxs],
        "mchars_per_s_e2e": round(sum(ix.n for ix in idxs) / min(walls3) / 1e3, 1)}
    for ix in idxs:
        ix.close()
    os.remove(path)
    os.rmdir(d)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default="gpurun_out/sweeps.json")
    ap.add_argument("--chars", type=int, default=100_000_000)
    ap.add_argument("--queries", type=int, default=1_000_000)
    ap.add_argument("--csv-mb", type=int, default=1024)
    ap.add_argument("--sections", default="c1,c5,c3")
    args = ap.parse_args()
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("sweeps.py needs a GPU")
    out = {"gpu": torch.cuda.get_device_name(0)}
    for sec in args.sections.split(","):
        t0 = time.perf_counter()
        {"c1": section_c1, "c3": section_c3, "c5": section_c5}[sec](args, out)
        out[f"{sec}_section_s"] = round(time.perf_counter() - t0, 1)
        with open(args.out, "w") as f:
            json.dump(out, f, indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
