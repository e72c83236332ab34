#!/usr/bin/env python
"""Generates tests/golden/*.npz by RUNNING THE REFERENCE (engine.c compiled unmodified into
oracle/_ref/libsa_ref.so from /root/reference).  Only runs in the authoring container; the
vectors are committed so the GPU box (no /root/reference) can check against them.

    python tests/golden/make_golden.py

Each case: text (u8), L, sa (u32, construct_truncated_suffix_array, engine.c:837), queries
(packed), start/end (get_substring_positions, engine.c:869) with `skipped` marking queries the
shim refuses because the reference would read out of bounds (mid == 0 wrap).
CSV case: file bytes, column, sa (file offsets) and per-rank quoted flags
(construct_truncated_suffix_array_from_csv_partitioned_mmap_full, engine.c:1454).
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import random_text, safe_queries_for  # noqa: E402
from oracle import pyoracle as O  # noqa: E402
from suffixarray_b200 import workloads as W  # noqa: E402


def text_case(name, text, L, rng, nq=200):
    text = np.ascontiguousarray(text, dtype=np.uint8)
    sa = O.ref_build(text, L)
    qs = safe_queries_for(text, rng, nq, max_len=48)
    s, e, skipped = O.ref_query(text, sa, L, qs)
    qbytes, qoff = O.pack_queries(qs)
    np.savez_compressed(os.path.join(HERE, name + ".npz"), text=text, L=np.uint32(L), sa=sa, qbytes=qbytes,
                        qoff=qoff, start=s, end=e, skipped=skipped)
    print(name, "n", text.size, "L", L, "queries", len(qs), "skipped", int(skipped.sum()))


def main():
    assert O.have_ref(), "needs /root/reference (authoring container)"
    rng = np.random.default_rng(20261104)
    text_case("banana", np.frombuffer(b"banana\nbandana", np.uint8), 32, rng, nq=20)
    names, _ = W.company_names_joined(1500, seed=5, upper=False)
    text_case("names_L32", names, 32, rng)
    text_case("names_L8", names, 8, rng)
    text_case("names_L64", names, 64, rng)
    eng = W.english_documents(30000, seed=6)
    text_case("english_L32", eng, 32, rng)
    text_case("english_L17", eng, 17, rng)
    text_case("sigma2_L32", random_text(rng, 20000, 2), 32, rng)
    text_case("sigma2_L16", random_text(rng, 20000, 2), 16, rng)
    text_case("sigma1_L64", random_text(rng, 3000, 1, newline_p=0.01), 64, rng)
    text_case("bytes_L32", np.concatenate([[10], rng.integers(1, 256, 8000)]).astype(np.uint8), 32, rng, nq=50)
    text_case("tiny_n33_L4", random_text(rng, 33, 3), 4, rng, nq=20)
    text_case("tiny_n32_L32", random_text(rng, 32, 3), 32, rng, nq=20)
    # CSV path
    data = W.csv_file_bytes(400, seed=44, quoted_fraction=0.25)
    with tempfile.NamedTemporaryFile(suffix=".csv", delete=False) as f:
        f.write(data)
        path = f.name
    try:
        out = {"csv_bytes": np.frombuffer(data, np.uint8)}
        for col in ("name", "city"):
            sa, q = O.ref_csv_build(path, col, 32, cap=len(data) + 16)
            out["sa_" + col] = sa
            out["quoted_" + col] = q
        np.savez_compressed(os.path.join(HERE, "csv_L32.npz"), **out)
        print("csv_L32", len(data), "bytes", {k: v.size for k, v in out.items()})
    finally:
        os.unlink(path)


if __name__ == "__main__":
    main()
