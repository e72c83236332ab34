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


def cpu_arm(text, qbytes, qoff, lower, L, threads=0, record_sample=2000):
    """The reference's engine.c (oracle/_ref) on the host cores beside a GPU number: build of the whole text,
    get_substring_positions for the whole batch (omp parallel-for), get_matching_records for a sample."""
    import bench
    cb = bench.cpu_reference_times(text, qbytes, qoff, lower, L, threads=threads, record_sample=record_sample)
    return {"kind": cb["kind"], "cores": cb["cores"], "build_s": round(cb["build_s"], 3),
            "build_mchars_per_s": round(cb["build_mchars_s"], 2), "queries": cb["query_sample"],
            "queries_per_s": round(cb["queries_per_s"], 1),
            "record_queries_per_s": round(cb["record_queries_per_s"], 1) if cb["record_queries_per_s"] else None}


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
        from oracle import pyoracle as O
        cores = O.ref_use_all_cores()
        tp = np.zeros(n + O.REF_PAD, dtype=np.uint8)
        tp[:n] = text
        ref_sa = np.empty(n, dtype=np.uint32)
        t0 = time.perf_counter()
        O.ref_build_padded(tp, n, L, ref_sa)
        cpu_s = time.perf_counter() - t0
        same = bool(np.array_equal(idx.sa(), ref_sa))
        del tp
        rows.append({"max_suffix_length": L, "build_ms": round(st["ms_total"], 3), "wall_ms": round(ms, 3),
                     "mchars_per_s": round(n / st["ms_total"] / 1e3, 1), "refine_rounds": st["rounds"],
                     "radix_passes": st["radix_passes"], "finish_ms": round(st["ms_finish"], 3),
                     "cpu_engine_c": {"cores": cores, "build_s": round(cpu_s, 3), "mchars_per_s": round(n / cpu_s / 1e6, 2)},
                     "sa_equal_to_engine_c": same})
        if L == 32:
            ref_sa32 = ref_sa
        if L != 32:
            idx.close()
        else:
            idx32 = idx
    out["c5_L_sweep"] = {"workload": f"{n} chars synthetic English (C2 text)", "rows": rows}
    # query-length sweep against the L=32 index, device-timed kernel (library CUDA events) + host-buffer wall time
    sa = idx32.sa()
    tpq = np.zeros(n + O.REF_PAD, dtype=np.uint8)
    tpq[:n] = text
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
        # CPU arm: get_substring_positions over the same batch on the same array, all cores
        import ctypes as C
        lens = (qoff[1:] - qoff[:-1]).astype(np.int64)
        zoff = np.zeros(nq + 1, dtype=np.uint64)
        np.cumsum(lens + 1, out=zoff[1:].view(np.int64))
        zb = np.zeros(int(zoff[nq]), dtype=np.uint8)
        zb[np.arange(int(qoff[nq]), dtype=np.int64) + np.repeat(np.arange(nq, dtype=np.int64), lens)] = qbytes[: int(qoff[nq])]
        rs = np.full(nq, 0xFFFFFFFF, np.uint32)
        re_ = np.full(nq, 0xFFFFFFFF, np.uint32)
        sk = np.zeros(nq, np.uint8)
        u32p, u64p, u8p = C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_uint8)
        t0 = time.perf_counter()
        O.ref_lib().ref_query_batch(tpq.ctypes.data, ref_sa32.ctypes.data_as(u32p), n, 32, zb.ctypes.data, zoff.ctypes.data_as(u64p),
                                    nq, rs.ctypes.data_as(u32p), re_.ctypes.data_as(u32p), sk.ctypes.data_as(u8p), cores)
        cpu_q = time.perf_counter() - t0
        keep = sk == 0
        same_q = bool(np.array_equal(start[keep], rs[keep]) and np.array_equal(end[keep], re_[keep]))
        qrows.append({"query_bytes": m, "kernel_ms": round(st["ms_query"], 4),
                      "cpu_engine_c_queries_per_s": round(nq / cpu_q, 0), "cpu_cores": cores, "ranges_equal_to_engine_c": same_q,
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
    # the drop-in class itself (tests/test.py style: query_records per name, and the batched entry)
    from suffixarray_b200 import SuffixArray
    docs = [d.decode() for d in text.tobytes().split(b"\n")]
    t0 = time.perf_counter()
    sa_cls = SuffixArray(documents=docs, max_suffix_length=32)
    cls_build_s = time.perf_counter() - t0
    qstr = [q.decode() for q in qs]
    t0 = time.perf_counter()
    for q1 in qstr[:1000]:
        sa_cls.query_records(q1, k=1000)
    one_us = (time.perf_counter() - t0) / 1000 * 1e6
    t0 = time.perf_counter()
    sa_cls.query_records_batch(qstr, k=1000)
    batch_s = time.perf_counter() - t0
    sa_cls.close()
    out["c1"]["class"] = {"construct_s_incl_join_encode": round(cls_build_s, 3), "query_records_us_per_call": round(one_us, 1),
                          "query_records_batch_10k_s": round(batch_s, 3),
                          "query_records_batch_per_s": round(len(qstr) / batch_s, 0),
                          "note": "suffixarray_b200.SuffixArray(documents=...): query_records one name at a time (tests/test.py:80-85 "
                                  "style) and query_records_batch; records are cut on the device (sa_query_records)"}
    # CPU arms: engine.c all cores and 1 core; libsais 1 thread and omp
    from oracle import pyoracle as O
    qb_, qo_ = O.pack_queries([q.lower() for q in qs])
    out["c1"]["cpu_engine_c_all_cores"] = cpu_arm(text, qb_, qo_, True, 32)
    out["c1"]["cpu_engine_c_1_core"] = cpu_arm(text, qb_, qo_, True, 32, threads=1)
    low = O.oracle_lowercase(text)
    for thr, name in ((1, "cpu_libsais_1_thread"), (O.ref_use_all_cores(), "cpu_libsais_omp")):
        t0 = time.perf_counter()
        O.ref_libsais(low, threads=thr)
        dt = time.perf_counter() - t0
        out["c1"][name] = {"threads": thr, "seconds": round(dt, 3), "mchars_per_s": round(n / dt / 1e6, 2),
                           "note": "full suffix array (baseline B), not ORDER-E"}
    idx.close()


def section_c3(args, out):
    from suffixarray_b200 import _lib as S
    from suffixarray_b200 import workloads as W
    d = tempfile.mkdtemp(prefix="sa_c3_")
    path = os.path.join(d, "c3.csv")
    t0 = time.perf_counter()
    rows, fbytes = W.write_csv_file(path, args.csv_mb)
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
    # ---- the drop-in class on the same file: construct + query_records (file semantics, the reference's
    #      record cut) beside the reference's own construct_..._mmap_full + get_matching_records_file
    from suffixarray_b200 import SuffixArray
    from oracle import pyoracle as O
    t0 = time.perf_counter()
    cls = SuffixArray(csv_file=path, search_column="name", max_suffix_length=32, reference_records=True)
    cls_s = time.perf_counter() - t0
    qs_all = W.unpack_queries(qbytes, qoff)
    qsel = [q for q in qs_all[:20000] if b'"' not in q and b"," not in q and 3 <= len(q) <= 32][:2000]
    qstr = [q.decode("latin-1") for q in qsel]
    t0 = time.perf_counter()
    one = [cls.query_records(q1, k=10) for q1 in qstr[:500]]
    one_us = (time.perf_counter() - t0) / 500 * 1e6
    t0 = time.perf_counter()
    got = cls.query_records_batch(qstr, k=10)
    batch_s = time.perf_counter() - t0
    out["c3"]["class"] = {"construct_s": round(cls_s, 3), "query_records_k10_us_per_call": round(one_us, 1),
                          "query_records_batch_k10_per_s": round(len(qstr) / batch_s, 0), "queries": len(qstr)}
    if O.have_ref() and not args.no_cpu:
        cores = O.ref_use_all_cores()
        t0 = time.perf_counter()
        R = O.RefCsvIndex(path, "name", 32)
        ref_s = time.perf_counter() - t0
        t0 = time.perf_counter()
        ref_recs = [R.records(q1, 10) for q1 in qsel]
        ref_q_s = time.perf_counter() - t0
        import csv as _csv
        compared = equal = 0
        unparsable = 0
        # hits ON the first byte of a row at a file offset >= 4096 are undefined in the reference: its backward seek
        # starts on the previous row's '\n' (engine.c:1262) and returns -1 as uint32 (:1277), the record is read from
        # wherever the forward seek left the file position -> such queries are counted apart, not compared
        _, _, hoff_q, pos_q, row_q = cls._parts[0].query(qsel, 10, True, False)
        rstart = cls._parts[0].row_starts()
        at_row_start = (pos_q.astype(np.uint64) == rstart[row_q.astype(np.int64)]) & (pos_q >= 4096)
        undefined = [bool(at_row_start[int(hoff_q[i]):int(hoff_q[i + 1])].any()) for i in range(len(qsel))]
        for qi, (g, want) in enumerate(zip(got, ref_recs)):
            if want is None:
                continue
            if undefined[qi]:
                unparsable += 1
                continue
            lines = [w.decode("utf-8", "replace").rstrip("\r") for w in want]
            # a hit on the first byte of a row at a file offset >= 4096 makes the reference's backward seek start ON
            # the previous row's '\n' (engine.c:1262) and return -1 as uint32 (:1277): its record is then garbage
            if any("\n" in ln for ln in lines):
                unparsable += 1
                continue
            try:
                want_rows = [dict(zip(cls.columns, row)) for row in _csv.reader(lines)]
            except _csv.Error:
                unparsable += 1
                continue
            compared += 1
            equal += int(g == want_rows)
        sa_ref, _ = R.arrays()
        out["c3"]["cpu_engine_c"] = {
            "cores": cores, "construct_mmap_full_s": round(ref_s, 2), "mchars_per_s_e2e": round(n / ref_s / 1e6, 2),
            "get_matching_records_file_k10_queries_per_s": round(len(qsel) / ref_q_s, 1),
            "get_matching_records_file_k10_us_per_query": round(ref_q_s / len(qsel) * 1e6, 1),
            "note": "construct_truncated_suffix_array_from_csv_partitioned_mmap_full (engine.c:1454) end to end; "
                    "get_matching_records_file (engine.c:1326), single thread as in the reference"}
        out["c3"]["parity"] = {"sa_equal_to_engine_c": bool(np.array_equal(cls.suffix_array, sa_ref)),
                               "record_lists_compared": compared, "record_lists_equal": equal,
                               "reference_record_lists_garbled": unparsable}
        R.close()
    cls.close()
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
    ap.add_argument("--no-cpu", action="store_true", help="skip the reference's CSV build in c3 (about a minute of CPU)")
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
