"""Pins the oracle (oracle/sa_oracle.c) against the UNMODIFIED reference compiled into
oracle/_ref/libsa_ref.so (oracle/Makefile; built from /root/reference where it lies).  The
reference's own tests hold no golden vectors (tests/test.py has no asserts), so this side-by-side
run plus tests/golden/ is what pins parity."""
import os
import tempfile

import numpy as np
import pytest

from conftest import GRID_L, GRID_N, GRID_SIGMA, random_text, safe_queries_for
from oracle import pyoracle as O
from suffixarray_b200 import workloads as W

needs_ref = pytest.mark.skipif(not O.have_ref(), reason="oracle/_ref/libsa_ref.so not built (no /root/reference)")


def need_ref():
    if not O.have_ref():
        pytest.skip("oracle/_ref/libsa_ref.so not built (no /root/reference)")


@needs_ref
@pytest.mark.parametrize("sigma", GRID_SIGMA)
def test_build_grid(sigma):
    rng = np.random.default_rng(sigma)
    for n in GRID_N:
        for L in GRID_L:
            t = random_text(rng, n, sigma)
            assert np.array_equal(O.oracle_build(t, L), O.ref_build(t, L)), (n, sigma, L)


@needs_ref
def test_build_edge_texts():
    rng = np.random.default_rng(3)
    cases = [
        np.full(3000, ord("a"), np.uint8), np.full(300, 10, np.uint8), np.full(300, 255, np.uint8),
        np.tile(np.frombuffer(b"abcabcabd\n", np.uint8), 300),
        rng.integers(1, 256, 5000).astype(np.uint8),
        np.frombuffer("naïve café\nüber straße\n日本語\n".encode() * 30, np.uint8),
        np.frombuffer(b"\n\n\nab\n\n\nab\n\n" * 20, np.uint8),
    ]
    for t in cases:
        for L in (1, 8, 9, 16, 17, 32, 64):
            assert np.array_equal(O.oracle_build(t, L), O.ref_build(t, L)), L


@needs_ref
def test_build_company_names():
    text, _ = W.company_names_joined(50_000, seed=41, upper=False)
    for L in (8, 32, 64):
        assert np.array_equal(O.oracle_build(text, L), O.ref_build(text, L))


@needs_ref
@pytest.mark.parametrize("sigma,L", [(2, 8), (3, 16), (8, 32), (26, 32), (26, 64)])
def test_query_matches_reference(sigma, L):
    rng = np.random.default_rng(50 + sigma + L)
    for n in (40, 500, 20000):
        t = random_text(rng, n, sigma)
        t[0] = 10
        sa = O.ref_build(t, L)
        qs = safe_queries_for(t, rng, 300, max_len=70)
        s, e = O.oracle_query(t, sa, L, qs)
        rs, re_, skipped = O.ref_query(t, sa, L, qs)
        keep = skipped == 0
        assert keep.sum() > 200
        assert np.array_equal(s[keep], rs[keep]) and np.array_equal(e[keep], re_[keep])


@needs_ref
def test_csv_scan_and_remap_match_reference():
    data = W.csv_file_bytes(3000, seed=43, quoted_fraction=0.2)
    with tempfile.NamedTemporaryFile(suffix=".csv", delete=False) as f:
        f.write(data)
        path = f.name
    try:
        for col, name in enumerate(["id", "name", "city", "score"]):
            text, mp, qd = O.oracle_csv_scan(data, col, 4)
            sa = O.oracle_build(text, 32)
            sa_file, q_rank = O.oracle_csv_remap(sa, mp, qd)
            rsa, rq = O.ref_csv_build(path, name, 32, cap=len(data) + 16)
            assert rsa.size == text.size, name
            assert np.array_equal(sa_file, rsa), name
            assert np.array_equal(q_rank, rq), name
    finally:
        os.unlink(path)


def test_hits_are_first_k_in_rank_order():
    text = np.frombuffer(b"\nabab abab\nabab\nxx", np.uint8)
    sa = O.oracle_build(text, 32)
    s, e = O.oracle_query(text, sa, 32, [b"ab", b"zz", b"xx"])
    off, pos, doc = O.oracle_hits(text, sa, s, e, 3)
    assert off.tolist() == [0, 3, 3, 4]
    assert pos[:3].tolist() == sa[s[0]:s[0] + 3].tolist()
    assert doc[3] == 3


def test_known_answer_banana():
    # produced by the reference itself during the survey
    t = np.frombuffer(b"banana\nbandana", np.uint8)
    sa = O.oracle_build(t, 32)
    assert sa.tolist() == [6, 13, 5, 11, 3, 1, 8, 0, 7, 10, 12, 4, 2, 9]
    s, e = O.oracle_query(t, sa, 32, [b"ana", b"ban"])
    assert (s.tolist(), e.tolist()) == ([3, 7], [5, 8])


def test_newline_prefix_closed_form():
    """scripts/check_wide.py derives the expected array of an index with more than 2^31 entries from
    SA('\\n' * j + R) = [0 .. j-1] ++ (SA(R) + j): the '\\n' bucket is never refined (engine.c:771-773).
    Checked here at small sizes against the oracle and, when built, the unmodified reference."""
    R = W.english_documents(60_000, seed=4001)
    sa_r = O.oracle_build(R, 32).astype(np.int64)
    assert R[sa_r[0]] == 10
    for j in (1, 16, 33, 5000):
        T = np.concatenate([np.full(j, 10, np.uint8), R])
        want = np.concatenate([np.arange(j, dtype=np.int64), sa_r + j])
        assert np.array_equal(O.oracle_build(T, 32).astype(np.int64), want), j
        if O.have_ref():
            assert np.array_equal(O.ref_build(T, 32).astype(np.int64), want), j


HAND_WRITTEN_CSV = {
    "plain": b"a,b,c\nx1,y1,z1\nx2,y2,z2\n",
    "no_trailing_newline": b"a,b,c\nx1,y1,z1\nx2,y2,z2",
    "no_trailing_newline_quoted": b'a,b,c\nx1,y1,"z,1"\nx2,y2,"z2"',
    "ends_with_comma": b"a,b,c\nx1,y1,z1\nx2,y2,",
    "empty_cells": b"a,b,c\n,,\nx,,z\n,y,\n",
    "quoted": b'a,b,c\n"hello, world",y1,"z""1"\n"",y2,"""q"""\n"multi\nline",y3,z3\n',
    "upper_and_crlf": b'a,b,c\r\nHeLLo,"WORLD",Z\r\n"Fo""O",Bar,baz\r\n',
    "ragged_rows": b"a,b,c\nx1,y1\nx2,y2,z2,w2\nx3\n",
    "one_column": b"a\nfoo\nbar\n\nbaz\n",
    "only_header": b"a,b,c\n",
    "only_header_no_newline": b"a,b,c",
    "escaped_pairs_at_edges": b'a,b\n"""x""","y"""\n"""""",""""\n',
    "high_bytes": b"a,b\n\xff\xfe,caf\xc3\xa9\n\x80,\x7f\n",
    "multiline_cells": b'a,b,c\n"line one\nline two",y1,z1\nx2,"p\nq\nr",z2\nx3,y3,"tail\ncell"\nx4,y4,z4\n',
    "crlf_multiline": b'a,b,c\r\nHeLLo,"WOR\r\nLD",Z\r\nfoo,Bar,baz\r\n',
}


@needs_ref
@pytest.mark.parametrize("name", sorted(HAND_WRITTEN_CSV))
def test_csv_oracle_matches_reference_on_hand_written_files(name):
    """The files the GPU column scanner is checked on (tests/test_gpu_csv_scan.py REGULAR,
    tests/test_gpu_csv_columns.py FILES) through the UNMODIFIED reference: text length, remapped array
    and quoted bits of every column.  (The reference keeps the '\\r' of a CRLF header line in the last
    column's name, engine.c:40-44 -- it is asked for that name; libsa_cuda strips it.)"""
    data = HAND_WRITTEN_CSV[name]
    first = data.split(b"\n")[0]
    header = first.rstrip(b"\r").decode().split(",")
    with tempfile.NamedTemporaryFile(suffix=".csv", delete=False) as f:
        f.write(data)
        path = f.name
    try:
        for col, cname in enumerate(header):
            text, mp, qd = O.oracle_csv_scan(data, col, len(header))
            sa_file, q_rank = O.oracle_csv_remap(O.oracle_build(text, 32), mp, qd)
            ref_name = cname + "\r" if (first.endswith(b"\r") and col == len(header) - 1) else cname
            rsa, rq = O.ref_csv_build(path, ref_name, 32, cap=len(data) + 64)
            assert rsa.size == text.size, cname
            assert np.array_equal(sa_file, rsa) and np.array_equal(q_rank, rq), cname
    finally:
        os.unlink(path)


@needs_ref
def test_csv_oracle_matches_reference_on_random_files():
    """60 random RFC-4180 files (1-4 columns, cells over a small alphabet that includes , " and raw
    newlines, written by Python's csv module; every third one without a trailing newline): the oracle's
    scan + build + remap equals the unmodified reference for every column."""
    import csv
    import io
    rng = np.random.default_rng(123)
    alpha = list('abcXYZ ,"\n.')
    for trial in range(60):
        ncols = int(rng.integers(1, 5))
        header = [f"c{i}" for i in range(ncols)]
        rows = [header] + [["".join(rng.choice(alpha, size=int(rng.integers(0, 9)))) for _ in range(ncols)]
                           for _ in range(int(rng.integers(1, 40)))]
        buf = io.StringIO()
        csv.writer(buf, lineterminator="\n").writerows(rows)
        data = buf.getvalue().encode()
        if trial % 3 == 0:
            data = data[:-1]
        with tempfile.NamedTemporaryFile(suffix=".csv", delete=False) as f:
            f.write(data)
            path = f.name
        try:
            for col, cname in enumerate(header):
                text, mp, qd = O.oracle_csv_scan(data, col, ncols)
                sa_file, q_rank = O.oracle_csv_remap(O.oracle_build(text, 32), mp, qd)
                rsa, rq = O.ref_csv_build(path, cname, 32, cap=len(data) + 64)
                assert rsa.size == text.size, (trial, cname)
                assert np.array_equal(sa_file, rsa) and np.array_equal(q_rank, rq), (trial, cname)
        finally:
            os.unlink(path)


# ----------------------------------------------------------------------------- CSV query path (row a9)
def _csv_cases():
    from suffixarray_b200 import workloads as W
    yield "quoted20", W.csv_file_bytes(3000, seed=11, quoted_fraction=0.2), "name", 1, 4
    yield "plain", W.csv_file_bytes(1500, seed=12, quoted_fraction=0.0), "city", 2, 4
    # a short file and a long max_suffix_length: probes cross the end of the file (stale line tail)
    yield "tiny", b"name,x\nabcabcabc,1\nabcab,2\nbcabc,3\nabc,4", "name", 0, 2


def _csv_queries(text, rng, nq):
    qs = []
    for _ in range(nq):
        a = int(rng.integers(0, max(1, text.size - 1)))
        m = int(rng.integers(1, 40))
        q = bytes(text[a:a + m]).split(b"\n")[0]
        if q and b"\0" not in q:
            qs.append(q)
    return qs + [b"inc", b"llc co", b"a b", b"corp,spring", b'co",lake', b"zzzzzz", b"7", b"abc", b"abcabcabc,1", b"c,4"]


@pytest.mark.parametrize("L", [8, 32, 64])
def test_file_query_oracle_vs_reference(L, tmp_path):
    """sao_query_file_batch == get_substring_positions_file (engine.c:920-999) on the reference's own
    CSV index, for every query the shim does not fence."""
    need_ref()
    rng = np.random.default_rng(300 + L)
    for name, data, column, col, ncols in _csv_cases():
        p = str(tmp_path / f"{name}.csv")
        with open(p, "wb") as f:
            f.write(data)
        R = O.RefCsvIndex(p, column, L)
        sa, _ = R.arrays()
        text, _, _ = O.oracle_csv_scan(data, col, ncols)
        qs = _csv_queries(text, rng, 600)
        rs, re_, sk = R.query(qs)
        ws, we = O.oracle_query_file(data, sa, L, qs)
        keep = sk == 0
        assert keep.sum() > len(qs) // 2, name
        assert np.array_equal(rs[keep], ws[keep]) and np.array_equal(re_[keep], we[keep]), name
        R.close()


def test_file_records_formula_vs_reference(tmp_path):
    """What SA_RECORDS_REFERENCE_COMPAT reproduces: get_matching_records_file (engine.c:1326-1390)
    returns file[row_start + (offset >= 4096) : row_end - 1] for a hit at file offset `offset` in the row
    [row_start, row_end) of the file (first byte dropped for offsets >= 4096, last byte always)."""
    need_ref()
    from suffixarray_b200 import workloads as W
    data = W.csv_file_bytes(2000, seed=13, quoted_fraction=0.1)
    p = str(tmp_path / "r.csv")
    with open(p, "wb") as f:
        f.write(data)
    R = O.RefCsvIndex(p, "name", 32)
    sa, _ = R.arrays()
    rows = O.oracle_csv_row_starts(data)
    text, _, _ = O.oracle_csv_scan(data, 1, 4)
    rng = np.random.default_rng(5)
    qs = [q for q in _csv_queries(text, rng, 300) if not q.startswith(b'"')]
    checked = 0
    for q in qs:
        recs = R.records(q, 7)
        if not recs:
            continue
        s, e, sk = R.query([q])
        assert sk[0] == 0
        take = min(7, int(e[0]) - int(s[0]) + 1)
        assert len(recs) == take
        for j in range(take):
            off = int(sa[int(s[0]) + j])
            r = int(np.searchsorted(rows[:-1], off, side="right") - 1)
            a, b = int(rows[r]), int(rows[r + 1]) - 1          # [a, b): the line without its '\n'
            want = data[a + (1 if off >= 4096 else 0): max(b - 1, a)]
            assert recs[j] == want, (q, j, off, recs[j], want)
            checked += 1
    assert checked > 200
    R.close()


def test_dedupe_oracle_small():
    off = np.array([0, 5, 5, 8], dtype=np.uint64)
    pos = np.array([10, 11, 12, 13, 14, 20, 21, 22], dtype=np.uint32)
    doc = np.array([3, 1, 3, 1, 2, 9, 9, 9], dtype=np.uint32)
    o, p, d = O.oracle_dedupe(off, pos, doc)
    assert o.tolist() == [0, 3, 3, 4] and p.tolist() == [10, 11, 14, 20] and d.tolist() == [3, 1, 2, 9]
