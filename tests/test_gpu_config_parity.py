"""Parity AT THE SIZE OF THE CONFIGS (BASELINE.json configs C1, C2 and a C3-shaped column), bit for bit
against the UNMODIFIED reference (oracle/_ref/libsa_ref.so = engine.c compiled where it lies):

* C1: 1 M company names (upper-cased input, lower-cased on the device), 10 k whole-name queries
* C2: the 100 M-char English document list, L = 32, the full 1 M-query batch
* C3-shaped: a CSV file whose search column holds >= 256 M chars, through sa_build_csv vs
  construct_truncated_suffix_array_from_csv_partitioned_mmap_full (engine.c:1454)

The reference builds these in seconds on the box's host cores; the GPU arrays must be identical:
`sa_copy_sa_u32` == `ref_build`, `sa_query_ranges` == `get_substring_positions` for every query the
reference can answer without its out-of-bounds wrap (ref_shim.c `skipped`), hit lists == the oracle's
hit expansion on the reference's ranges."""
import os
import tempfile

import numpy as np
import pytest

from oracle import pyoracle as O
from suffixarray_b200 import _lib as S
from suffixarray_b200 import workloads as W

pytestmark = pytest.mark.gpu

L = 32


def _need_ref():
    if not O.have_ref():
        pytest.skip("oracle/_ref/libsa_ref.so not present")


def _zero_terminated(qbytes, qoff, nq):
    lens = (qoff[1:nq + 1] - qoff[:nq]).astype(np.int64)
    zoff = np.zeros(nq + 1, dtype=np.uint64)
    np.cumsum(lens + 1, out=zoff[1:].view(np.int64))
    zbytes = np.zeros(int(zoff[nq]), dtype=np.uint8)
    src = qbytes[: int(qoff[nq])]
    dst = np.arange(src.size, dtype=np.int64) + np.repeat(np.arange(nq, dtype=np.int64), lens)
    zbytes[dst] = src
    return zbytes, zoff


def _ref_ranges(tp, n, sa, qbytes, qoff):
    import ctypes as C
    nq = int(qoff.size - 1)
    zbytes, zoff = _zero_terminated(qbytes, qoff, nq)
    start = np.full(nq, 0xFFFFFFFF, dtype=np.uint32)
    end = np.full(nq, 0xFFFFFFFF, dtype=np.uint32)
    skipped = np.zeros(nq, dtype=np.uint8)
    u32p, u64p, u8p = C.POINTER(C.c_uint32), C.POINTER(C.c_uint64), C.POINTER(C.c_uint8)
    O.ref_lib().ref_query_batch(tp.ctypes.data, sa.ctypes.data_as(u32p), n, L, zbytes.ctypes.data,
                                zoff.ctypes.data_as(u64p), nq, start.ctypes.data_as(u32p), end.ctypes.data_as(u32p),
                                skipped.ctypes.data_as(u8p), len(os.sched_getaffinity(0)))
    return start, end, skipped


def _check_text_config(text, qbytes, qoff, lowercase, hit_sample):
    _need_ref()
    low = O.oracle_lowercase(text) if lowercase else text
    n = int(low.size)
    tp = np.zeros(n + O.REF_PAD, dtype=np.uint8)
    tp[:n] = low
    want = np.empty(n, dtype=np.uint32)
    O.ref_build_padded(tp, n, L, want)                       # the reference itself, all host cores

    idx = S.Index(L)
    idx.build_text(text, lowercase=lowercase)
    got = idx.sa()
    if not np.array_equal(got, want):
        bad = np.flatnonzero(got != want)
        raise AssertionError(f"SA differs from engine.c at {bad.size} of {n} ranks, first {bad[:5]}")
    del got

    rs, re_, skipped = _ref_ranges(tp, n, want, qbytes, qoff)
    keep = skipped == 0
    assert keep.mean() > 0.99
    s, e, off = idx.query_ranges_packed(qbytes, qoff, k=1000)
    assert np.array_equal(s[keep], rs[keep]), f"start differs at {np.flatnonzero(keep & (s != rs))[:5]}"
    assert np.array_equal(e[keep], re_[keep]), f"end differs at {np.flatnonzero(keep & (e != re_))[:5]}"
    # hit lists of a sample of the batch (oracle expansion of the reference's own ranges)
    pos, doc = idx.query_hits(int(off[-1]))
    rng = np.random.default_rng(7)
    pick = np.sort(rng.choice(np.flatnonzero(keep), size=min(hit_sample, int(keep.sum())), replace=False))
    woff, wpos, wdoc = O.oracle_hits(low, want, rs[pick], re_[pick], 1000)
    for t, q in enumerate(pick.tolist()):
        a, b = int(off[q]), int(off[q + 1])
        wa, wb = int(woff[t]), int(woff[t + 1])
        assert b - a == wb - wa, q
        assert np.array_equal(pos[a:b], wpos[wa:wb]) and np.array_equal(doc[a:b], wdoc[wa:wb]), q
    idx.close()


def test_c1_full_names():
    text, starts = W.company_names_joined(1_000_000, seed=1001, upper=True)
    qs = W.whole_name_queries(text, starts, 10_000, seed=1002)
    qbytes, qoff = O.pack_queries([q.lower() for q in qs])
    _check_text_config(text, qbytes, qoff, True, 10_000)


def test_c2_full_english_100m():
    text = W.english_documents(100_000_000, seed=2001)
    qbytes, qoff = W.substring_queries(text, 1_000_000, seed=2002)
    _check_text_config(text, qbytes, qoff, False, 20_000)


def test_c3_shaped_csv_column_256m():
    _need_ref()
    d = tempfile.mkdtemp(prefix="sa_c3p_")
    path = os.path.join(d, "c3.csv")
    try:
        rows, fbytes = W.write_csv_file(path, 256, seed=3001, quoted_rows_per_chunk=20_000)
        assert fbytes < (1 << 31)
        want_sa, want_q = O.ref_csv_build(path, "name", L, cap=fbytes + 16)
        assert want_sa.size >= 256_000_000
        idx = S.Index(L)
        idx.build_csv(path, "name")
        assert idx.n == want_sa.size
        got = idx.sa()
        if not np.array_equal(got, want_sa):
            bad = np.flatnonzero(got != want_sa)
            raise AssertionError(f"CSV SA differs from engine.c at {bad.size} ranks, first {bad[:5]}")
        assert np.array_equal(idx.quoted_bits(), want_q)
        idx.close()
    finally:
        try:
            os.unlink(path)
            os.rmdir(d)
        except OSError:
            pass
