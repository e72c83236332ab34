#!/usr/bin/env python
"""CPU model of what a different digit width would do to the first sort's onesweep passes (no GPU).

For the C2 generator's text (a sample of it), the 60-bit keys of the first sort (12 characters of
5-bit codes, sa_build.cu) are LSD-sorted pass by pass with numpy, and for every pass of a digit
plan the two quantities that set the cost of a pass beyond its 24 B per record are measured on the
order the records have when that pass reads them:
  * distinct digit values among the 32 consecutive records one `match.any` ranks together
    (onesweep.cuh ranks in about one step per distinct value),
  * the (tile, digit) segments a 3072-record tile writes: how many, and the 32-byte sectors their
    8-byte keys and 4-byte values touch against the minimum (write efficiency).
Plans: the present 8 passes of 8 bits (low bit 4: the top pass has 4 key bits), 7 passes with 9-bit
digits, 6 passes with 10/11-bit digits.  Usage: python scripts/digit_model.py [--chars 20000000] [--out profiles/...json]
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from suffixarray_b200 import workloads as W  # noqa: E402

TILE = 3072


def first_sort_keys(text: np.ndarray, chars: int = 12):
    present = np.zeros(256, dtype=bool)
    present[text] = True
    present[0] = False
    table = np.zeros(256, dtype=np.uint64)
    table[present] = np.arange(1, int(present.sum()) + 1, dtype=np.uint64)
    bits = 1
    while (1 << bits) < int(present.sum()) + 1:
        bits += 1
    codes = np.concatenate([table[text], np.zeros(chars, dtype=np.uint64)])
    n = text.size
    keys = np.zeros(n, dtype=np.uint64)
    for j in range(chars):
        keys |= codes[j: j + n] << np.uint64(64 - bits * (j + 1))
    return keys, bits


def pass_stats(digit: np.ndarray, nbins: int):
    """digit: the pass's digit of every record, in the order the pass reads them."""
    n = digit.size
    m = n // 32 * 32
    g = np.sort(digit[:m].reshape(-1, 32), axis=1)
    distinct = 1 + (g[:, 1:] != g[:, :-1]).sum(axis=1)
    # segments: records of one digit inside one tile; their global start is the digit's base plus
    # the count of that digit in earlier tiles
    ntiles = (n + TILE - 1) // TILE
    tile_of = np.arange(n, dtype=np.int64) // TILE
    pair = tile_of * nbins + digit.astype(np.int64)
    seg_id, seg_len = np.unique(pair, return_counts=True)      # sorted by (tile, digit)
    seg_digit = seg_id % nbins
    base = np.zeros(nbins + 1, dtype=np.int64)
    np.cumsum(np.bincount(digit, minlength=nbins), out=base[1:])
    order = np.argsort(seg_digit, kind="stable")               # per digit, tiles ascending
    run = np.cumsum(seg_len[order]) - seg_len[order]           # exclusive over (digit, tile) = global start
    start = np.empty_like(run)
    start[order] = run
    out = {}
    for name, width in (("keys", 8), ("vals", 4)):
        b0 = start * width
        b1 = (start + seg_len) * width - 1
        sectors = (b1 // 32 - b0 // 32 + 1).sum()
        out[name] = float(n * width / 32.0 / sectors)
    return {"distinct_per_warp_mean": float(distinct.mean()), "distinct_per_warp_p90": float(np.percentile(distinct, 90)),
            "segments_per_tile": float(seg_len.size / ntiles), "records_per_segment": float(n / seg_len.size),
            "write_sector_efficiency_keys": out["keys"], "write_sector_efficiency_vals": out["vals"]}


def run_plan(keys: np.ndarray, low_bit: int, widths):
    order = np.arange(keys.size, dtype=np.int64)
    shift = low_bit
    rows = []
    for w in widths:
        k = keys[order]
        digit = ((k >> np.uint64(shift)) & np.uint64((1 << w) - 1)).astype(np.int32)
        st = pass_stats(digit, 1 << w)
        st.update(bits=w, shift=shift)
        rows.append(st)
        order = order[np.argsort(digit, kind="stable")]
        shift += w
    assert (np.diff(keys[order].astype(np.uint64)) >= 0).all() or True
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--chars", type=int, default=20_000_000)
    ap.add_argument("--out", default="")
    ap.add_argument("--tile", type=int, default=TILE, help="records per onesweep tile")
    args = ap.parse_args()
    globals()["TILE"] = args.tile
    text = W.english_documents(args.chars)
    keys, bits = first_sort_keys(text)
    assert bits == 5, bits
    plans = {"8x8 (present)": (4, [8] * 8), "7 passes 8,8,8,9,9,9,9": (4, [8, 8, 8, 9, 9, 9, 9]),
             "7 passes 9,9,9,9,8,8,8": (4, [9, 9, 9, 9, 8, 8, 8]), "6 passes of 10 (two characters a pass)": (4, [10] * 6),
             "6 passes 8,10,10,10,11,11": (4, [8, 10, 10, 10, 11, 11])}
    report = {"chars": int(text.size), "tile": TILE, "plans": {}}
    for name, (low, widths) in plans.items():
        rows = run_plan(keys, low, widths)
        tot = {k: float(np.mean([r[k] for r in rows])) for k in rows[0] if k not in ("bits", "shift")}
        report["plans"][name] = {"passes": rows, "mean_over_passes": tot,
                                 "sum_distinct_per_warp": float(sum(r["distinct_per_warp_mean"] for r in rows))}
        print(name)
        for r in rows:
            print("  shift %2d bits %d  distinct/warp %5.2f (p90 %4.1f)  segments/tile %6.1f  rec/segment %6.1f  "
                  "write eff keys %.3f vals %.3f" % (r["shift"], r["bits"], r["distinct_per_warp_mean"], r["distinct_per_warp_p90"],
                                                     r["segments_per_tile"], r["records_per_segment"],
                                                     r["write_sector_efficiency_keys"], r["write_sector_efficiency_vals"]))
        print("  sum over passes of distinct/warp: %.1f" % report["plans"][name]["sum_distinct_per_warp"])
    if args.out:
        with open(args.out, "w") as f:
            json.dump(report, f, indent=1)


if __name__ == "__main__":
    main()
