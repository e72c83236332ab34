"""The Cython host side (suffixarray_b200/suffix_array_cy.pyx: the reference's own language for this
layer, `cdef extern from "sa_cuda.h"` in place of `engine.h`) through the same drop-in checks as the
ctypes mirror: README snippets, ValueError / RuntimeError behaviour, hit lists in the oracle's order,
the 2 GiB partition walk, CSV rows (multi-line cells), several search columns, save / load."""
import csv
import io
import os
import tempfile

import numpy as np
import pytest

from oracle import pyoracle as O
from suffixarray_b200 import SuffixArray as PySuffixArray
from suffixarray_b200 import workloads as W

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def CySuffixArray():
    try:                                           # the in-tree extension travels with the snapshot
        from suffixarray_b200.suffix_array_cy import SuffixArray
    except ImportError:
        from suffixarray_b200 import _lib
        _lib.build_cython_extension()
        from suffixarray_b200.suffix_array_cy import SuffixArray
    return SuffixArray


def test_readme_documents_snippet(CySuffixArray):
    docs = ["Hello World", "hello there", "Goodbye", "say hello", ""]
    sa = CySuffixArray(documents=docs, max_suffix_length=32)
    assert sorted(sa.query_records("hello", k=10)) == sorted(["hello world", "hello there", "say hello"])
    assert sa.query_records("") == []
    assert sa.query_records("zebra") == []
    assert sa.query_records("HELLO T") == ["hello there"]          # query is lower-cased (pyx:228)
    sa.close()


def test_head_methods_and_errors(CySuffixArray):
    sa = CySuffixArray(max_suffix_length=16)
    with pytest.raises(ValueError):
        sa.construct_truncated_suffix_array_documents(12345)       # pyx:130-134
    with pytest.raises(RuntimeError):
        sa.query_records("x")
    sa.construct_truncated_suffix_array_documents(("ab", "cab", "b"))
    assert sorted(sa.query_records("ab")) == ["ab", "cab"]
    with pytest.raises(ValueError):
        CySuffixArray(documents=["a"], csv_file="x.csv")
    sa.close()


def test_documents_match_oracle_and_ctypes_mirror(CySuffixArray):
    names = W.company_names(3000, seed=77, upper=True)
    docs = [b.decode() for b in names]
    sa = CySuffixArray(documents=docs, max_suffix_length=32)
    text = O.oracle_lowercase(np.frombuffer("\n".join(docs).encode(), np.uint8))
    want_sa = O.oracle_build(text, 32)
    assert np.array_equal(sa.suffix_array, want_sa)
    qs = ["inc", "group", docs[5].lower(), docs[17][:4].lower(), "qqqq", ""]
    s, e = O.oracle_query(text, want_sa, 32, [q.encode() for q in qs[:-1]])
    off, pos, doc = O.oracle_hits(text, want_sa, s, e, 25)
    got = sa.query_records_batch(qs, k=25)
    low_docs = [d.lower() for d in docs]
    for i, q in enumerate(qs[:-1]):
        assert got[i] == [low_docs[d] for d in doc[int(off[i]):int(off[i + 1])]]
    assert got[-1] == []
    py = PySuffixArray(documents=docs, max_suffix_length=32)
    assert got == py.query_records_batch(qs, k=25)
    py.close()
    sa.close()


def test_partitioned_documents_walk(CySuffixArray):
    names = W.company_names(400, seed=91, upper=True)
    docs = [b.decode() for b in names]
    sa = CySuffixArray(max_suffix_length=32)
    sa.PARTITION_BYTES = 1500                                       # pyx:148-152 cuts every 2 GiB
    sa.construct_truncated_suffix_array_documents(docs)
    text = O.oracle_lowercase(np.frombuffer("\n".join(docs).encode(), np.uint8))
    arrays = sa.suffix_array
    assert len(arrays) == -(-text.size // 1500) and len(arrays) > 3
    for p, got in enumerate(arrays):               # partition p is sorted on `text + p * cut` of the ONE joined buffer (pyx:176-180)
        n_p = min(1500, text.size - p * 1500)
        assert np.array_equal(got, O.oracle_build_tail(text[p * 1500: p * 1500 + n_p + 960], n_p, 32))
    low = [d.lower() for d in docs]
    starts = np.concatenate([[0], np.flatnonzero(text == 10) + 1])
    for q, k in (("inc", 7), ("a", 40), ("group", 1000), (low[13], 5)):
        want = []
        for p in range(len(arrays)):                                # the reference's walk, on the oracle
            part = text[p * 1500:(p + 1) * 1500]
            s, e = O.oracle_query_tail(text[p * 1500: p * 1500 + part.size + 960], part.size, arrays[p], 32, [q.encode()])
            off, pos, _ = O.oracle_hits(part, arrays[p], s, e, k - len(want))
            want += [low[int(np.searchsorted(starts, int(x) + p * 1500, side="right") - 1)] for x in pos]
            if len(want) >= k:
                break
        assert sa.query_records(q, k=k) == want
    sa.close()


def test_csv_rows_columns_and_save_load(CySuffixArray, tmp_path):
    rows = [["id", "name", "city", "note"]]
    for i in range(300):
        name = f"Acme {i} Holdings" if i % 3 else f"Globex\nUnit {i}"        # raw newline inside quotes
        city = "Springfield" if i % 2 else "Shelbyville, East"
        rows.append([str(i), name, city, f"acme note {i}" if i % 50 == 0 else "n/a"])
    buf = io.StringIO()
    csv.writer(buf, lineterminator="\n").writerows(rows)
    path = str(tmp_path / "multi.csv")
    with open(path, "wb") as f:
        f.write(buf.getvalue().encode())
    dict_rows = [dict(zip(rows[0], r)) for r in rows]

    one = CySuffixArray(csv_file=path, search_column="name", max_suffix_length=32, csv_query="text")
    assert one.columns == rows[0]
    assert one.query_records("unit 297", k=10) == [dict_rows[298]]
    assert one.query_records("name")[0] == dict(zip(rows[0], rows[0]))     # header is row 0
    py = PySuffixArray(csv_file=path, search_column="name", max_suffix_length=32, csv_query="text")
    assert np.array_equal(one.suffix_array, py.suffix_array)
    for q in ("acme", "globex", "9 hold", "zzz"):
        assert one.query_records(q, k=40) == py.query_records(q, k=40)
    py.close()
    one.close()
    with pytest.raises(Exception):
        CySuffixArray(csv_file=path, search_column="nope")
    with pytest.raises(ValueError):
        CySuffixArray(csv_file=path)

    sa = CySuffixArray(csv_file=path, search_column=["name", "note", "city"], max_suffix_length=32, csv_query="text")
    assert sa.search_columns == ["name", "note", "city"] and len(sa.suffix_array) == 3
    name_hits = sa.query_records("acme", k=1000, column="name")
    note_hits = sa.query_records("acme", k=1000, column="note")
    assert len(name_hits) == 200 and len(note_hits) == 6
    assert sa.query_records("acme", k=1000) == name_hits + note_hits
    assert sa.query_records("acme", k=203) == name_hits + note_hits[:3]
    with pytest.raises(ValueError):
        sa.query_records("x", column="id")
    p = str(tmp_path / "multi.sa")
    sa.save(p)
    sb = CySuffixArray.load(p)
    sc = PySuffixArray.load(p)                                      # the two host sides share the container
    for q in ("acme", "unit 297", "east", "zzz"):
        assert sa.query_records(q, k=300) == sb.query_records(q, k=300) == sc.query_records(q, k=300)
    for x in (sa, sb, sc):
        x.close()
