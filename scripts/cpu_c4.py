#!/usr/bin/env python
"""CPU baselines for BASELINE config C4 (one 4 G-char corpus), timed on the GPU box's host cores:

  A  the reference's live path the way it would handle such a text (suffix_array.pyx:148-180): sequential
     partitions of (just under) 2 GiB, construct_truncated_suffix_array (engine.c:837) on each, all cores.
     Partition size is 2^31 - 4096 instead of 2^31: at exactly 2^31 engine.c:738 (`int char_idx =
     suffix_array[i] + current_depth`) overflows for the last positions of a partition.
  B  libsais64_omp (vendored libsais64.c, /root/reference/suffix_array/libsais64.h:86), all cores: a FULL
     int64 suffix array of the whole text -- timing baseline only (BASELINE.json "libsais64 / int64 SA path").

    python scripts/cpu_c4.py --chars 4000000000 --out gpurun_out/r2_cpu_c4.json
The text is the C4 bench text: 8 x (C2 generator at 500 M chars, seeds as bench.py uses for ranks 0..7)."""
import argparse
import json
import os
import sys
import time
from multiprocessing import Pool

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def gen(args):
    rank, chars = args
    from suffixarray_b200 import workloads as W
    CH = 250_000_000
    parts, left, k = [], chars, 0
    while left > 0:
        take = min(CH, left) - (1 if left > CH else 0)
        parts.append(W.english_documents(take, seed=2001 + rank + 1000 * k))
        left -= take
        if left > 0:
            parts.append(np.frombuffer(b"\n", np.uint8))
            left -= 1
        k += 1
    return np.concatenate(parts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--chars", type=int, default=4_000_000_000)
    ap.add_argument("--ranks", type=int, default=8)
    ap.add_argument("--out", default="gpurun_out/r2_cpu_c4.json")
    ap.add_argument("--skip-libsais64", action="store_true")
    a = ap.parse_args()
    from oracle import pyoracle as O
    cores = O.ref_use_all_cores()
    per = a.chars // a.ranks
    t0 = time.perf_counter()
    with Pool(min(a.ranks, cores)) as pool:
        pieces = pool.map(gen, [(r, per) for r in range(a.ranks)])
    n = sum(p.size for p in pieces) + a.ranks - 1
    buf = np.zeros(n + O.REF_PAD, dtype=np.uint8)
    at = 0
    for r, p in enumerate(pieces):
        buf[at:at + p.size] = p
        at += p.size
        if r < a.ranks - 1:
            buf[at] = 10
            at += 1
    del pieces
    gen_s = time.perf_counter() - t0
    out = {"n_chars": int(n), "cores": int(cores), "generate_s": round(gen_s, 1)}
    # ---- A: partitions of just under 2 GiB, one after the other
    PB = (1 << 31) - 4096
    parts = []
    tA = 0.0
    off = 0
    while off < n:
        m = min(PB, n - off)
        sa = np.empty(m, dtype=np.uint32)
        t0 = time.perf_counter()
        O.ref_build_padded(buf[off:], m, 32, sa)          # text + offset of the ONE buffer, like pyx:176-180
        dt = time.perf_counter() - t0
        parts.append({"offset": int(off), "n": int(m), "seconds": round(dt, 2)})
        tA += dt
        off += m
        del sa
    out["A_engine_c_partitions"] = {"partition_bytes": PB, "partitions": parts, "seconds": round(tA, 2),
                                    "mchars_per_s": round(n / tA / 1e6, 2),
                                    "note": "construct_truncated_suffix_array per partition, sequentially (suffix_array.pyx:148-180), "
                                            "max_suffix_length 32, all host cores"}
    print(json.dumps(out), flush=True)
    if not a.skip_libsais64:
        t0 = time.perf_counter()
        sa64 = O.ref_libsais64(buf[:n], threads=cores)
        tB = time.perf_counter() - t0
        out["B_libsais64_omp"] = {"seconds": round(tB, 2), "mchars_per_s": round(n / tB / 1e6, 2),
                                  "sa_dtype": "int64", "first_entries": sa64[:4].tolist(),
                                  "note": "full (untruncated) suffix array of the whole text; timing baseline only"}
        del sa64
    os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
    with open(a.out, "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
