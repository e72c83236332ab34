"""GPU parity for the query side of the boundary beyond get_substring_positions:

* SA_QUERY_FILE_SEMANTICS == get_substring_positions_file (engine.c:920-999), the bisection behind the
  reference's only Python-reachable query (suffix_array.pyx:221-247): vs the oracle restatement and vs
  the UNMODIFIED reference on its own CSV index;
* hit lists of a CSV index as rows of the file; the row table (sa_copy_row_starts);
* SA_QUERY_DEDUPE (segmented dedupe by record id; the reference keeps duplicates, engine.c:1187);
* sa_query_records (records gathered on the device) incl. SA_RECORDS_REFERENCE_COMPAT ==
  get_matching_records_file's record strings (engine.c:1326-1390);
* per-call result objects: two threads querying one index at the same time.
"""
import os
import threading

import numpy as np
import pytest

from oracle import pyoracle as O
from suffixarray_b200 import _lib as S
from suffixarray_b200 import workloads as W

pytestmark = pytest.mark.gpu


def _cases():
    yield "quoted20", W.csv_file_bytes(3000, seed=11, quoted_fraction=0.2), "name", 1, 4
    yield "plain", W.csv_file_bytes(1500, seed=12, quoted_fraction=0.0), "city", 2, 4
    yield "tiny", b"name,x\nabcabcabc,1\nabcab,2\nbcabc,3\nabc,4", "name", 0, 2
    yield "big", W.csv_file_bytes(60_000, seed=14, quoted_fraction=0.05), "name", 1, 4


def _queries(text, rng, nq):
    qs = []
    for _ in range(nq):
        a = int(rng.integers(0, max(1, text.size - 1)))
        m = int(rng.integers(1, 40))
        q = bytes(text[a:a + m]).split(b"\n")[0]
        if q and b"\0" not in q:
            qs.append(q)
    return qs + [b"inc", b"llc co", b"a b", b"corp,spring", b'co",lake', b"zzzzzz", b"7", b"abc", b"abcabcabc,1", b"c,4",
                 b"x" * 100]


def _write(tmp_path, name, data):
    p = str(tmp_path / f"{name}.csv")
    with open(p, "wb") as f:
        f.write(data)
    return p


@pytest.mark.parametrize("L", [8, 32, 64])
def test_file_semantics_vs_oracle_and_reference(L, tmp_path):
    rng = np.random.default_rng(700 + L)
    for name, data, column, col, ncols in _cases():
        p = _write(tmp_path, name, data)
        idx = S.Index(L)
        idx.build_csv(p, column)
        sa_file = idx.sa()
        text, _, _ = O.oracle_csv_scan(data, col, ncols)
        qs = _queries(text, rng, 2000 if name == "big" else 500)
        res = idx.result()
        s, e, off = res.ranges(qs, k=50, flags=S.SA_QUERY_FILE_SEMANTICS)
        ws, we = O.oracle_query_file(data, sa_file, L, qs)
        assert np.array_equal(s, ws), (name, np.flatnonzero(s != ws)[:5])
        assert np.array_equal(e, we), (name, np.flatnonzero(e != we)[:5])
        if O.have_ref():
            R = O.RefCsvIndex(p, column, L)
            rsa, _ = R.arrays()
            assert np.array_equal(rsa, sa_file)
            rs, re_, sk = R.query(qs)
            keep = sk == 0
            assert keep.sum() > len(qs) // 2
            assert np.array_equal(s[keep], rs[keep]) and np.array_equal(e[keep], re_[keep]), name
            R.close()
        # hits: file offsets + rows of the file
        rows = O.oracle_csv_row_starts(data)
        woff, wpos, wdoc = O.oracle_hits_rows(sa_file, rows, ws, we, 50)
        hoff, pos, doc = res.hits()
        assert np.array_equal(hoff, woff) and np.array_equal(off, woff)
        assert np.array_equal(pos, wpos) and np.array_equal(doc, wdoc), name
        assert np.array_equal(idx.row_starts(), rows), name
        # the in-memory semantics still answer through the same result object
        s2, e2, _ = res.ranges(qs, k=50)
        sa_text = O.oracle_build(text, L)
        ms, me = O.oracle_query(text, sa_text, L, qs)
        assert np.array_equal(s2, ms) and np.array_equal(e2, me), name
        res.close()
        idx.close()


def test_file_semantics_multi_column_and_errors(tmp_path):
    data = W.csv_file_bytes(2000, seed=21, quoted_fraction=0.1)
    p = _write(tmp_path, "m", data)
    idxs = [S.Index(32), S.Index(32)]
    S.Index.build_csv_columns(idxs, p, ["name", "city"])
    for ix, col in zip(idxs, (1, 2)):
        text, _, _ = O.oracle_csv_scan(data, col, 4)
        rng = np.random.default_rng(col)
        qs = _queries(text, rng, 300)
        res = ix.result()
        s, e, _ = res.ranges(qs, k=10, flags=S.SA_QUERY_FILE_SEMANTICS)
        ws, we = O.oracle_query_file(data, ix.sa(), 32, qs)
        assert np.array_equal(s, ws) and np.array_equal(e, we), col
        res.close()
    for ix in idxs:
        ix.close()
    t = S.Index(32)
    t.build_text(b"\nabc\nabd")
    res = t.result()
    with pytest.raises(S.SaCudaError):
        res.ranges([b"abc"], flags=S.SA_QUERY_FILE_SEMANTICS)      # not a CSV index
    res.close()
    t.close()


@pytest.mark.parametrize("k", [7, 1000, 3000])
def test_dedupe_text_index(k):
    # few long documents with repeated words: many hits share a record
    rng = np.random.default_rng(k)
    words = [b"alpha", b"beta", b"gamma", b"delta", b"abab", b"ab"]
    docs = [b" ".join(words[int(i)] for i in rng.integers(0, len(words), size=int(rng.integers(1, 60)))) for _ in range(400)]
    text = np.frombuffer(b"\n" + b"\n".join(docs), np.uint8)
    idx = S.Index(32)
    idx.build_text(text)
    sa = O.oracle_build(text, 32)
    qs = [b"alpha", b"ab", b"a", b"gamma delta", b"zz", b"beta", b" "]
    ws, we = O.oracle_query(text, sa, 32, qs)
    woff, wpos, wdoc = O.oracle_hits(text, sa, ws, we, k)
    doff, dpos, ddoc = O.oracle_dedupe(woff, wpos, wdoc)
    res = idx.result()
    res.ranges(qs, k=k)
    off, pos, doc = res.hits(dedupe=True)
    assert np.array_equal(off, doff) and np.array_equal(pos, dpos) and np.array_equal(doc, ddoc)
    assert int(off[-1]) <= int(woff[-1]) and (k < 1000 or int(off[-1]) < int(woff[-1]))      # something was removed
    # without the flag the lists are untouched
    off2, pos2, doc2 = res.hits()
    assert np.array_equal(off2, woff) and np.array_equal(pos2, wpos) and np.array_equal(doc2, wdoc)
    # records of the de-duplicated lists = the documents themselves
    hoff, roff, rb, rdoc = res.records(dedupe=True)
    assert np.array_equal(hoff, doff) and np.array_equal(rdoc, ddoc)
    b = rb.tobytes()
    alldocs = [b""] + docs
    for j in range(rdoc.size):
        assert b[int(roff[j]):int(roff[j + 1])] == alldocs[int(rdoc[j])], j
    res.close()
    idx.close()


def test_records_csv_default_and_reference_compat(tmp_path):
    data = W.csv_file_bytes(2000, seed=13, quoted_fraction=0.1)
    p = _write(tmp_path, "r", data)
    idx = S.Index(32)
    idx.build_csv(p, "name")
    sa_file = idx.sa()
    rows = O.oracle_csv_row_starts(data)
    text, _, _ = O.oracle_csv_scan(data, 1, 4)
    rng = np.random.default_rng(5)
    qs = [q for q in _queries(text, rng, 300) if not q.startswith(b'"')]
    res = idx.result()
    s, e, off = res.ranges(qs, k=7, flags=S.SA_QUERY_FILE_SEMANTICS)
    hoff, roff, rb, rdoc = res.records()
    b = rb.tobytes()
    assert np.array_equal(hoff, off)
    for j in range(rdoc.size):                                   # default: the whole line without its '\n'
        r = int(rdoc[j])
        assert b[int(roff[j]):int(roff[j + 1])] == data[int(rows[r]): int(rows[r + 1]) - 1], j
    # dedupe on rows
    doff, droff, drb, ddoc = res.records(dedupe=True)
    _, wpos, wdoc = O.oracle_hits_rows(sa_file, rows, s, e, 7)
    woff2, _, wdoc2 = O.oracle_dedupe(off, wpos, wdoc)
    assert np.array_equal(doff, woff2) and np.array_equal(ddoc, wdoc2)
    if O.have_ref():
        R = O.RefCsvIndex(p, "name", 32)
        hoff, roff, rb, rdoc = res.records(compat=True)
        b = rb.tobytes()
        checked = 0
        for i, q in enumerate(qs):
            want = R.records(q, 7)
            if want is None:
                continue
            a, z = int(hoff[i]), int(hoff[i + 1])
            got = [b[int(roff[j]):int(roff[j + 1])] for j in range(a, z)]
            assert got == want, (q, got[:2], want[:2])
            checked += len(want)
        assert checked > 200
        R.close()
    res.close()
    idx.close()


def test_two_threads_query_one_index():
    text = W.english_documents(2_000_000, seed=41)
    idx = S.Index(32)
    idx.build_text(text)
    sa = idx.sa()
    batches = []
    for t in range(2):
        qb, qo = W.substring_queries(text, 30_000, seed=50 + t)
        qs = W.unpack_queries(qb, qo)
        ws, we = O.oracle_query(text, sa, 32, qs)
        woff, wpos, wdoc = O.oracle_hits(text, sa, ws, we, 100)
        batches.append((qb, qo, ws, we, woff, wpos, wdoc))
    errors = []

    def worker(t):
        try:
            qb, qo, ws, we, woff, wpos, wdoc = batches[t]
            res = idx.result()
            for _ in range(5):
                s, e, off = res.ranges_packed(qb, qo, k=100)
                hoff, pos, doc = res.hits()
                assert np.array_equal(s, ws) and np.array_equal(e, we)
                assert np.array_equal(off, woff) and np.array_equal(pos, wpos) and np.array_equal(doc, wdoc)
            res.close()
        except Exception as ex:           # noqa: BLE001
            errors.append(repr(ex))

    th = [threading.Thread(target=worker, args=(t,)) for t in range(2)]
    for x in th:
        x.start()
    for x in th:
        x.join()
    assert not errors, errors
    idx.close()


def test_attach_csv_file_after_load(tmp_path):
    data = W.csv_file_bytes(1500, seed=31, quoted_fraction=0.1)
    p = _write(tmp_path, "a", data)
    idx = S.Index(32)
    idx.build_csv(p, "name")
    qs = [b"inc", b"llc", b"corp", b"a"]
    res = idx.result()
    want = res.ranges(qs, k=5, flags=S.SA_QUERY_FILE_SEMANTICS)
    want_rec = res.records()
    res.close()
    saved = str(tmp_path / "a.idx")
    idx.save(saved)
    idx.close()
    idx2 = S.Index(32)
    idx2.load(saved)
    res = idx2.result()
    with pytest.raises(S.SaCudaError):
        res.ranges(qs, k=5, flags=S.SA_QUERY_FILE_SEMANTICS)       # the file is not on the device yet
    idx2.attach_csv_file(p)
    got = res.ranges(qs, k=5, flags=S.SA_QUERY_FILE_SEMANTICS)
    for a, b in zip(want, got):
        assert np.array_equal(a, b)
    got_rec = res.records()
    for a, b in zip(want_rec, got_rec):
        assert np.array_equal(a, b)
    res.close()
    idx2.close()
    os.unlink(saved)


def test_partition_look_ahead_tail():
    """sa_build_text_tail: a partition of a longer text is sorted -- and queried -- on `text + cut` of the
    ONE buffer (suffix_array.pyx:176-180), like the reference, not on a zero-padded slice."""
    names = W.company_names(400, seed=91, upper=False)
    text = np.frombuffer(b"\n".join(names), np.uint8)
    pb = 1500
    tp = np.zeros(text.size + O.REF_PAD, np.uint8)
    tp[: text.size] = text
    differs = 0
    for p in range(-(-text.size // pb)):
        n = min(pb, text.size - p * pb)
        tail = min(960, text.size - p * pb - n)
        chunk = np.ascontiguousarray(text[p * pb: p * pb + n + tail])
        idx = S.Index(32)
        idx.build_text(chunk, tail=tail)
        got = idx.sa()
        want = O.oracle_build_tail(chunk, n, 32)
        assert np.array_equal(got, want), p
        if O.have_ref():
            ref = np.empty(n, np.uint32)
            O.ref_build_padded(tp[p * pb:], n, 32, ref)
            assert np.array_equal(got, ref), p
        differs += int(not np.array_equal(want, O.oracle_build(chunk[:n], 32)))
        qs = [b"inc", b"a", b"group", bytes(chunk[n - 3: n + 4]).split(b"\n")[0] or b"x"]
        s, e, _ = idx.query_ranges(qs, k=10)
        ws, we = O.oracle_query_tail(chunk, n, want, 32, qs)
        assert np.array_equal(s, ws) and np.array_equal(e, we), p
        idx.close()
    assert differs > 0            # the look-ahead matters on this text


def test_class_csv_defaults_to_reference_semantics(tmp_path):
    """SuffixArray(csv_file=...) answers like the reference's query_records: file-semantics ranges
    (get_substring_positions_file) and, with reference_records=True, its record strings."""
    from suffixarray_b200 import SuffixArray
    data = W.csv_file_bytes(2000, seed=13, quoted_fraction=0.1)
    p = _write(tmp_path, "c", data)
    sa = SuffixArray(csv_file=p, search_column="name", max_suffix_length=32, reference_records=True)
    sa_file = sa.suffix_array
    rows = O.oracle_csv_row_starts(data)
    text, _, _ = O.oracle_csv_scan(data, 1, 4)
    rng = np.random.default_rng(9)
    qs = [q for q in _queries(text, rng, 200) if not q.startswith(b'"') and b"," not in q]
    ws, we = O.oracle_query_file(data, sa_file, 32, qs)
    woff, wpos, wdoc = O.oracle_hits_rows(sa_file, rows, ws, we, 5)
    import csv
    got = sa.query_records_batch([q.decode() for q in qs], k=5)
    for i, q in enumerate(qs):
        lines = []
        for h in range(int(woff[i]), int(woff[i + 1])):
            r, off = int(wdoc[h]), int(wpos[h])
            a, b = int(rows[r]), int(rows[r + 1]) - 1
            lines.append(data[a + (1 if off >= 4096 else 0): max(b - 1, a)].decode().rstrip("\r"))
        want = [dict(zip(sa.columns, row)) for row in csv.reader(lines)]
        assert got[i] == want, q
    sa.close()
    # "text" semantics + whole rows + dedupe: every row at most once, and it really contains the query
    sb = SuffixArray(csv_file=p, search_column="name", max_suffix_length=32, csv_query="text", dedupe=True)
    for q in ("inc", "a", "llc"):
        recs = sb.query_records(q, k=50)
        assert recs and all(q in r["name"].lower() for r in recs)
        ids = [r["id"] for r in recs]
        assert len(ids) == len(set(ids))
    sb.close()
