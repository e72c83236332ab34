"""Host logic of the drop-in class (`suffixarray_b200/_host.py`, the mirror of suffix_array.pyx:110-267) on CPU:
the class body is shared by the Cython and the ctypes binding and only talks to a small backend interface, so
here it runs over a stand-in backend that answers from the ORACLE (test infrastructure; the product bindings
reach libsa_cuda.so and nothing else).  Covered: join / encode / lower-casing of queries, the 2 GiB partition
walk with look-ahead tails (pyx:148-180, :221-247), decoding of the packed record buffer, k, empty queries,
dedupe across partitions, the search-column walk of a CSV index, argument errors (pyx:130-134)."""
import numpy as np
import pytest

from oracle import pyoracle as O
from suffixarray_b200._host import SuffixArrayHost


class OracleIndex:
    """Backend stand-in: ORDER-E, bisection and hit expansion from oracle/sa_oracle.c."""
    builds = []          # (bytes handed over, tail) of every build_text call, in order

    def __init__(self, max_suffix_length=64, device=0):
        self.L, self.n, self._t, self._sa = int(max_suffix_length), 0, None, None

    def build_text(self, arr, lowercase=False, tail=0):
        a = np.ascontiguousarray(arr, dtype=np.uint8)
        OracleIndex.builds.append((a.tobytes(), int(tail)))
        self._t = O.oracle_lowercase(a) if lowercase else a.copy()
        self.n = int(a.size - tail)
        self._sa = O.oracle_build_tail(self._t, self.n, self.L)

    def sa(self):
        return self._sa

    def text(self):
        return self._t[: self.n]

    def close(self):
        pass

    def _hits(self, queries, k, dedupe):
        s, e = O.oracle_query_tail(self._t, self.n, self._sa, self.L, queries)
        off, pos, doc = O.oracle_hits(self._t[: self.n], self._sa, s, e, k)
        if dedupe:
            off, pos, doc = O.oracle_dedupe(off, pos, doc)
        return s, e, off, pos, doc

    def query(self, queries, k, file_semantics, dedupe):
        assert not file_semantics
        return self._hits(queries, k, dedupe)

    def query_records(self, queries, k, file_semantics, dedupe, compat):
        _, _, off, _, doc = self._hits(queries, k, dedupe)
        t = self._t[: self.n].tobytes()
        starts = [0] + [i + 1 for i, b in enumerate(t) if b == 10]
        recs = []
        for d in doc.tolist():
            a = starts[d]
            b = starts[d + 1] - 1 if d + 1 < len(starts) else len(t)
            recs.append(t[a:b])
        roff = np.zeros(len(recs) + 1, dtype=np.uint64)
        np.cumsum([len(r) for r in recs], out=roff[1:].view(np.int64))
        return off, roff, np.frombuffer(b"".join(recs), dtype=np.uint8), doc


class CpuSuffixArray(SuffixArrayHost):
    _Index = OracleIndex


DOCS = ["The quick brown fox", "jumps over the lazy dog", "THE END", "quick quick quick", "a fox, a dog, the lot",
        "nothing here", "Fox and Hound", "the the the"] * 6


def test_documents_single_partition():
    sa = CpuSuffixArray(max_suffix_length=32, documents=DOCS)
    assert len(sa._parts) == 1
    for q in ("fox", "THE", "quick", "zzz", "dog", "lot"):
        got = sa.query_records(q, k=1000)
        assert all(q.lower() in r for r in got), q
        assert all(r == r.lower() for r in got)                       # records are the lower-cased documents
        assert len(got) == sum(d.lower().count(q.lower()) for d in DOCS), q      # one record per hit, duplicates kept
    assert sa.query_records("", k=10) == []                           # pyx:210-211
    assert len(sa.query_records("the", k=5)) == 5
    batch = sa.query_records_batch(["fox", "", "dog"], k=3)
    assert [len(b) for b in batch] == [3, 0, 3]
    # dedupe: every document at most once per query
    dd = CpuSuffixArray(max_suffix_length=32, documents=DOCS, dedupe=True)
    got = dd.query_records("quick", k=1000)
    assert len(got) == 12 and set(got) == {"quick quick quick", "the quick brown fox"}   # 2 documents x 6 copies: 12 ids, 24 hits
    assert len(sa.query_records("quick", k=1000)) == 24
    sa.close()
    dd.close()


def test_documents_argument_handling():
    sa = CpuSuffixArray(max_suffix_length=16, documents=(d for d in DOCS[:4]))       # any iterable (pyx:130-134)
    assert sa.query_records("fox", k=10) == ["the quick brown fox"]
    with pytest.raises(ValueError):
        CpuSuffixArray(max_suffix_length=16, documents=12345)
    with pytest.raises(ValueError):
        CpuSuffixArray(documents=DOCS, csv_file="x.csv")
    with pytest.raises(ValueError):
        CpuSuffixArray(csv_file="x.csv")                              # needs search_column=
    with pytest.raises(ValueError):
        CpuSuffixArray(csv_query="neither")
    with pytest.raises(RuntimeError):
        CpuSuffixArray().query_records("a")
    sa.close()


def test_partition_walk_hands_over_look_ahead_tails():
    class Small(CpuSuffixArray):
        PARTITION_BYTES = 300
    OracleIndex.builds = []
    sa = Small(max_suffix_length=32, documents=DOCS)
    joined = "\n".join(DOCS).encode()
    pb = 300
    nparts = -(-len(joined) // pb)
    assert len(sa._parts) == nparts and len(OracleIndex.builds) == nparts
    for i, (data, tail) in enumerate(OracleIndex.builds):             # pyx:173-180: cut at fixed offsets, successor visible
        end = min(len(joined), (i + 1) * pb)
        assert tail == min(960, len(joined) - end)
        assert data == joined[i * pb: end + tail]
    whole = CpuSuffixArray(max_suffix_length=32, documents=DOCS)
    for q in ("fox", "the", "quick", "dog", "hound", "nothing here"):
        got = sa.query_records(q, k=1000)
        want = whole.query_records(q, k=1000)
        assert sorted(got) == sorted(want), q                         # same hits, found partition by partition
        assert all(q in r for r in got)
        # documents that straddle a cut come back whole
        assert all(r in [d.lower() for d in DOCS] for r in got)
    assert len(sa.query_records("the", k=4)) == 4                     # the walk stops when k records are found
    dd = Small(max_suffix_length=32, documents=DOCS, dedupe=True)
    got = dd.query_records("the", k=1000)
    assert len(got) == len({(r, i) for i, r in enumerate(got)})       # list of documents, each index once
    docs_with = [d for d in DOCS if "the" in d.lower()]
    assert len(got) == len(docs_with)
    sa.close(); whole.close(); dd.close()


class StubCsvIndex:
    """A CSV index that returns canned rows: exercises the column walk and the row decoding only."""
    rows = {}

    def __init__(self, max_suffix_length=64, device=0):
        self.column = None
        self.calls = []

    def build_csv(self, path, column):
        self.column = column

    def close(self):
        pass

    def query_records(self, queries, k, file_semantics, dedupe, compat):
        self.calls.append((list(queries), k, file_semantics, dedupe, compat))
        recs, off = [], [0]
        for q in queries:
            mine = StubCsvIndex.rows.get((self.column, q), [])[:k]
            recs += mine
            off.append(len(recs))
        roff = np.zeros(len(recs) + 1, dtype=np.uint64)
        np.cumsum([len(r) for r in recs], out=roff[1:].view(np.int64))
        return (np.asarray(off, dtype=np.uint64), roff, np.frombuffer(b"".join(recs), dtype=np.uint8),
                np.arange(len(recs), dtype=np.uint32))


def _stub_build_csv_columns(parts, path, cols):
    for p, c in zip(parts, cols):
        p.build_csv(path, c)


class CsvHost(SuffixArrayHost):
    _Index = StubCsvIndex
    _build_csv_columns = staticmethod(_stub_build_csv_columns)


def test_csv_column_walk_and_row_decoding(tmp_path):
    p = tmp_path / "t.csv"
    p.write_bytes(b'id,name,city\r\n1,"Acme, Inc",Paris\r\n2,Bolt,Rome\r\n')
    StubCsvIndex.rows = {("name", b"acme"): [b'1,"Acme, Inc",Paris\r'], ("city", b"acme"): [b"9,Zed,Acme City"],
                         ("city", b"rome"): [b"2,Bolt,Rome", b"3,Cog,Rome"]}
    sa = CsvHost(max_suffix_length=32, csv_file=str(p), search_column=["name", "city"], reference_records=True)
    assert sa.columns == ["id", "name", "city"] and sa.search_columns == ["name", "city"]
    got = sa.query_records("ACME", k=5)                               # queries are lower-cased (pyx:228); both columns walked
    assert got == [{"id": "1", "name": "Acme, Inc", "city": "Paris"}, {"id": "9", "name": "Zed", "city": "Acme City"}]
    assert sa._parts[0].calls[-1] == ([b"acme"], 5, True, False, True)         # file semantics + reference cut passed on
    assert sa._parts[1].calls[-1][1] == 4                             # the second column is asked for what is still missing
    assert sa.query_records("acme", k=1) == got[:1]
    assert len(sa._parts[1].calls) == 1                               # k reached in the first column: the walk stopped before the second
    assert sa.query_records("rome", k=5, column="city") == [{"id": "2", "name": "Bolt", "city": "Rome"},
                                                            {"id": "3", "name": "Cog", "city": "Rome"}]
    with pytest.raises(ValueError):
        sa.query_records("rome", column="id")
    with pytest.raises(ValueError):
        CsvHost(csv_file=str(p), search_column=["name", "name"])
    with pytest.raises(FileNotFoundError):
        CsvHost(csv_file=str(tmp_path / "missing.csv"), search_column="name")
    text = CsvHost(max_suffix_length=32, csv_file=str(p), search_column="name", csv_query="text")
    text.query_records("acme")
    assert text._parts[0].calls[-1][2] is False                       # csv_query="text": in-memory semantics
    sa.close(); text.close()
