"""SURVEY 8(f)-4 and the row ids of CSV hits.

* sa_build_csv_columns: several columns of one file in one call (one read / upload / quote analysis)
  must give, per column, exactly what the oracle's restatement of the reference's single-column path
  gives (engine.c:461-654): column text, ORDER-E, file offsets, quoted bits.
* sa_query_hits on a CSV index: hit_pos = file offset (the remapped array, engine.c:647-651),
  hit_doc = row of the FILE holding that offset (record ends = '\\n' outside quotes, header = row 0)
  -- the record the reference seeks from the offset (engine.c:1217-1323) -- also for cells with raw
  newlines inside quotes and for ragged rows, where counting '\\n' bytes of the column text drifts.
"""
import csv
import io
import os
import tempfile

import numpy as np
import pytest

from oracle import pyoracle as O
from suffixarray_b200 import SuffixArray
from suffixarray_b200 import workloads as W

pytestmark = pytest.mark.gpu

L = 32


def expected(data: bytes, col: int, ncols: int):
    text, mp, qd = O.oracle_csv_scan(data, col, ncols)
    sa = O.oracle_build(text, L)
    sa_file, q_rank = O.oracle_csv_remap(sa, mp, qd)
    return text, sa, sa_file, q_rank


def file_rows(data: bytes, offsets: np.ndarray) -> np.ndarray:
    """row of the file that holds each offset: record ends ('\\n' outside quotes) in front of it"""
    b = np.frombuffer(data, np.uint8)
    inq = (np.cumsum(b == ord('"')) - (b == ord('"'))) & 1          # quotes in front of each byte
    ends = np.flatnonzero((b == 10) & (inq == 0))
    return np.searchsorted(ends, offsets.astype(np.int64), side="left")


def write_tmp(data: bytes) -> str:
    with tempfile.NamedTemporaryFile(suffix=".csv", delete=False) as f:
        f.write(data)
        return f.name


FILES = {
    "plain": (b"a,b,c\nx1,y1,z1\nx2,y2,z2\n", 2),
    "multiline_cells": (b'a,b,c\n"line one\nline two",y1,z1\nx2,"p\nq\nr",z2\nx3,y3,"tail\ncell"\nx4,y4,z4\n', 2),
    "ragged_rows": (b"a,b,c\nx1,y1\nx2,y2,z2,w2\nx3\nx4,y4,z4\n", 2),
    "no_trailing_newline": (b'a,b,c\nx1,y1,"z,1"\nx2,y2,"z\n2"', 2),
    "crlf": (b'a,b,c\r\nHeLLo,"WOR\r\nLD",Z\r\nfoo,Bar,baz\r\n', 2),
    "irregular_quote": (b'a,b,c\nab"c,d"e,f\nx,"y\ny",z\nq,r,s\n', 0),       # host column scan, device row table
}


@pytest.mark.parametrize("name", sorted(FILES))
def test_columns_in_one_call_match_single_column_oracle(name):
    from suffixarray_b200 import _lib as S
    data, scanner = FILES[name]
    header = data.split(b"\n")[0].rstrip(b"\r").decode().split(",")
    path = write_tmp(data)
    idxs = [S.Index(L) for _ in header]
    try:
        idxs[0].debug_set("csv_scanner", scanner)
        order = header[::-1]                                    # any order of columns
        S.Index.build_csv_columns(idxs, path, order)
        for idx, colname in zip(idxs, order):
            text, sa, sa_file, q_rank = expected(data, header.index(colname), len(header))
            assert idx.n == text.size, colname
            assert np.array_equal(idx.text(), text), colname
            assert np.array_equal(idx.text_positions(), sa), colname
            assert np.array_equal(idx.sa(), sa_file), colname
            assert np.array_equal(idx.quoted_bits(), q_rank), colname
            # every suffix of the column as a query: positions are file offsets, rows are file rows
            qs = sorted({bytes(text[p:p + 3]).split(b"\n")[0] for p in range(text.size)} - {b""})
            s, e, off = idx.query_ranges(qs, k=1000)
            pos, row = idx.query_hits(int(off[-1]))
            for i in range(len(qs)):
                if s[i] == S.UINT32_MAX:
                    continue
                want_pos = sa_file[int(s[i]):int(e[i]) + 1][:1000]
                assert np.array_equal(pos[int(off[i]):int(off[i + 1])], want_pos), (colname, qs[i])
            assert np.array_equal(row, file_rows(data, pos)), colname
        with pytest.raises(Exception):
            S.Index.build_csv_columns(idxs[:2], path, ["a", "nope"])
        with pytest.raises(Exception):
            S.Index.build_csv_columns([idxs[0], idxs[0]], path, ["a", "b"])
    finally:
        for idx in idxs:
            idx.close()
        os.unlink(path)


def test_generated_csv_columns_and_rows():
    from suffixarray_b200 import _lib as S
    data = W.csv_file_bytes(5000, seed=31, quoted_fraction=0.3)
    header = ["id", "name", "city", "score"]
    path = write_tmp(data)
    idxs = [S.Index(L) for _ in range(3)]
    single = S.Index(L)
    try:
        S.Index.build_csv_columns(idxs, path, ["name", "city", "id"], timing=True)
        for idx, colname in zip(idxs, ["name", "city", "id"]):
            text, sa, sa_file, q_rank = expected(data, header.index(colname), 4)
            assert np.array_equal(idx.sa(), sa_file) and np.array_equal(idx.quoted_bits(), q_rank)
            single.build_csv(path, colname)
            assert np.array_equal(single.sa(), idx.sa()) and np.array_equal(single.text(), idx.text())
            st = idx.stats()
            assert st["csv_host_scan"] == 0 and st["ms_total"] > 0
        rng = np.random.default_rng(5)
        text = idxs[0].text()
        qs = [bytes(text[p:p + int(rng.integers(2, 9))]).split(b"\n")[0] or b"a" for p in rng.integers(0, text.size - 9, 400)]
        s, e, off = idxs[0].query_ranges(qs, k=50)
        pos, row = idxs[0].query_hits(int(off[-1]))
        assert np.array_equal(row, file_rows(data, pos))
        assert row.max() <= data.count(b"\n")
    finally:
        for idx in idxs + [single]:
            idx.close()
        os.unlink(path)


def test_suffix_array_class_multi_column_and_multiline_rows(tmp_path):
    rows = [["id", "name", "city", "note"]]
    for i in range(300):
        name = f"Acme {i} Holdings" if i % 3 else f"Globex\nUnit {i}"        # raw newline inside quotes
        city = "Springfield" if i % 2 else "Shelbyville, East"
        rows.append([str(i), name, city, f"acme note {i}" if i % 50 == 0 else "n/a"])
    buf = io.StringIO()
    csv.writer(buf, lineterminator="\n").writerows(rows)
    data = buf.getvalue().encode()
    path = str(tmp_path / "multi.csv")
    with open(path, "wb") as f:
        f.write(data)
    dict_rows = [dict(zip(rows[0], r)) for r in rows]

    one = SuffixArray(csv_file=path, search_column="name", max_suffix_length=L, csv_query="text")
    got = one.query_records("unit 297", k=10)                    # after 99 multi-line cells: rows must not drift
    assert got == [dict_rows[298]] and got[0]["name"] == "Globex\nUnit 297"
    got = one.query_records("acme 100 ", k=10)
    assert got == [dict_rows[101]]
    one.close()

    sa = SuffixArray(csv_file=path, search_column=["name", "note", "city"], max_suffix_length=L, csv_query="text")
    assert sa.search_columns == ["name", "note", "city"] and len(sa.suffix_array) == 3
    name_hits = sa.query_records("acme", k=1000, column="name")
    note_hits = sa.query_records("acme", k=1000, column="note")
    assert len(name_hits) == 200 and all("acme" in r["name"].lower() for r in name_hits)
    assert len(note_hits) == 6 and all(r["note"].startswith("acme note") for r in note_hits)
    assert sa.query_records("acme", k=1000) == name_hits + note_hits     # columns in the order given
    assert sa.query_records("acme", k=150) == name_hits[:150]
    assert sa.query_records("acme", k=203) == name_hits + note_hits[:3]
    east = sa.query_records("ville, ea", k=5)
    assert len(east) == 5 and all(r["city"] == "Shelbyville, East" for r in east)
    with pytest.raises(ValueError):
        sa.query_records("x", column="id")
    batch = sa.query_records_batch(["acme", "", "springf"], k=4)
    assert batch[0] == name_hits[:4] and batch[1] == [] and len(batch[2]) == 4
    p = str(tmp_path / "multi.sa")
    sa.save(p)
    sb = SuffixArray.load(p)
    assert sb.search_columns == sa.search_columns
    for q in ("acme", "unit 297", "east", "zzz"):
        assert sa.query_records(q, k=300) == sb.query_records(q, k=300)
    sa.close()
    sb.close()


@pytest.mark.parametrize("world,exchange,replicate", [(2, "p2p", "fused"), (3, "p2p", "allgather"), (4, "nccl", "allgather")])
def test_sharded_csv_column_matches_single_gpu(world, exchange, replicate):
    """BASELINE config 3 over several GPUs: the sample-sort build fed by the CSV column scan
    (sa_shard_begin_csv), ranks emulated on one GPU; array of file offsets, quoted bits and hits must
    equal the single-GPU sa_build_csv result (itself checked against the oracle above)."""
    from suffixarray_b200 import _lib as S
    from suffixarray_b200 import sharded as SH
    cases = [(W.csv_file_bytes(3000, seed=61, quoted_fraction=0.3), "name", 0),
             (FILES["multiline_cells"][0], "a", 0),
             (FILES["irregular_quote"][0], "b", 0),
             (W.csv_file_bytes(800, seed=62, quoted_fraction=0.2), "city", 1)]
    for data, column, scanner in cases:
        path = write_tmp(data)
        single = S.Index(L)
        ranks = []
        try:
            single.debug_set("csv_scanner", scanner)
            single.build_csv(path, column)
            _, ranks = SH.build_emulated(None, L, world, exchange=exchange, replicate=replicate, csv=(path, column))
            text = single.text()
            qs = sorted({bytes(text[p:p + 4]).split(b"\n")[0] for p in range(0, text.size, 7)} - {b""})[:300]
            ws, we, woff = single.query_ranges(qs, k=20)
            wpos, wrow = single.query_hits(int(woff[-1]))
            for r in ranks:
                assert r.idx.n == single.n
                assert np.array_equal(r.idx.sa(), single.sa())
                assert np.array_equal(r.idx.text_positions(), single.text_positions())
                assert np.array_equal(r.idx.quoted_bits(), single.quoted_bits())
                s, e, off = r.idx.query_ranges(qs, k=20)
                pos, row = r.idx.query_hits(int(off[-1]))
                assert np.array_equal(s, ws) and np.array_equal(e, we) and np.array_equal(off, woff)
                assert np.array_equal(pos, wpos) and np.array_equal(row, wrow)
        finally:
            single.close()
            for r in ranks:
                r.close()
            os.unlink(path)
