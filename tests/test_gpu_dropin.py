"""Drop-in class tests: suffixarray_b200.SuffixArray mirrors suffix_array.pyx:110-267 and the
README snippets (README.md:14-28, :34-50); record lists are checked against the oracle's hit
lists and against brute-force substring search."""
import os
import tempfile

import numpy as np
import pytest

from oracle import pyoracle as O
from suffixarray_b200 import SuffixArray
from suffixarray_b200 import _lib as S
from suffixarray_b200 import workloads as W

pytestmark = pytest.mark.gpu


def test_readme_documents_snippet():
    docs = ["Hello World", "hello there", "Goodbye", "say hello", ""]
    sa = SuffixArray(documents=docs, max_suffix_length=32)
    got = sa.query_records("hello", k=10)
    assert sorted(got) == sorted(["hello world", "hello there", "say hello"])
    assert sa.query_records("") == []
    assert sa.query_records("zebra") == []
    assert sa.query_records("HELLO T") == ["hello there"]      # query is lower-cased (pyx:228)
    sa.close()


def test_head_methods_and_value_error():
    sa = SuffixArray(max_suffix_length=16)
    with pytest.raises(ValueError):
        sa.construct_truncated_suffix_array_documents(12345)   # pyx:130-134
    with pytest.raises(RuntimeError):
        sa.query_records("x")
    sa.construct_truncated_suffix_array_documents(("ab", "cab", "b"))   # non-list is list()-ed
    assert sorted(sa.query_records("ab")) == ["ab", "cab"]
    sa.close()


def test_documents_hits_match_oracle_order():
    names = W.company_names(3000, seed=77, upper=True)
    docs = [b.decode() for b in names]
    sa = SuffixArray(documents=docs, max_suffix_length=32)
    text = O.oracle_lowercase(np.frombuffer("\n".join(docs).encode(), np.uint8))
    want_sa = O.oracle_build(text, 32)
    assert np.array_equal(sa.suffix_array, want_sa)
    qs = ["inc", "group", docs[5].lower(), docs[17][:4].lower(), "qqqq"]
    s, e = O.oracle_query(text, want_sa, 32, [q.encode() for q in qs])
    off, pos, doc = O.oracle_hits(text, want_sa, s, e, 25)
    got = sa.query_records_batch(qs, k=25)
    low_docs = [d.lower() for d in docs]
    for i, q in enumerate(qs):
        want = [low_docs[d] for d in doc[int(off[i]):int(off[i + 1])]]
        assert got[i] == want
        assert all(q in r for r in got[i])
        n_true = sum(r.count(q) for r in low_docs)
        assert len(got[i]) == min(25, n_true)
    sa.close()


def test_csv_constructor_rows():
    data = W.csv_file_bytes(500, seed=9, quoted_fraction=0.3)
    with tempfile.NamedTemporaryFile(suffix=".csv", delete=False) as f:
        f.write(data)
        path = f.name
    try:
        sa = SuffixArray(csv_file=path, search_column="name", max_suffix_length=32, csv_query="text")
        assert sa.columns == ["id", "name", "city", "score"]
        import csv
        import io
        rows = list(csv.DictReader(io.StringIO(data.decode())))
        target = rows[123]["name"].lower()
        got = sa.query_records(target, k=50)
        assert any(r["id"] == rows[123]["id"] for r in got)
        assert all(target in r["name"].lower() for r in got)
        assert sa.query_records("name") and sa.query_records("name")[0] == dict(zip(sa.columns, sa.columns))  # header is row 0
        with pytest.raises(Exception):
            SuffixArray(csv_file=path, search_column="nope")
        sa.close()
    finally:
        os.unlink(path)


def test_partitioned_documents_walk(monkeypatch):
    """suffix_array.pyx:148-180 / :221-247: fixed-size byte partitions, each with its own suffix array
    (bit-exact vs the oracle per partition); a query walks them in order until k records are found."""
    names = W.company_names(400, seed=91, upper=True)
    docs = [b.decode() for b in names]
    monkeypatch.setattr(SuffixArray, "PARTITION_BYTES", 1500)
    sa = SuffixArray(documents=docs, max_suffix_length=32)
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
        for p in range(len(arrays)):                        # the reference's walk, on the oracle
            part = text[p * 1500:(p + 1) * 1500]
            s, e = O.oracle_query_tail(text[p * 1500: p * 1500 + part.size + 960], part.size, arrays[p], 32, [q.encode()])
            off, pos, _ = O.oracle_hits(part, arrays[p], s, e, k - len(want))
            want += [low[int(np.searchsorted(starts, int(x) + p * 1500, side="right") - 1)] for x in pos]
            if len(want) >= k:
                break
        assert sa.query_records(q, k=k) == want
    sa.close()


def test_save_load_round_trip(tmp_path):
    """a saved index answers queries after load without being rebuilt; the
    reference-layout export holds (u64 start, u64 end, u32 L, u32 n, u32 SA[n])."""
    from suffixarray_b200 import _lib as S
    names = W.company_names(3000, seed=17, upper=True)
    docs = [b.decode() for b in names]
    sa = SuffixArray(documents=docs, max_suffix_length=32)
    p = str(tmp_path / "docs.sa")
    sa.save(p)
    sb = SuffixArray.load(p)
    assert np.array_equal(sa.suffix_array, sb.suffix_array)
    for q in ("inc", "group", docs[5].lower(), "zzzz"):
        assert sa.query_records(q, k=50) == sb.query_records(q, k=50)
    sa._idx.export_reference_files(p + ".ref", p + ".refq")
    raw = np.fromfile(p + ".ref", dtype=np.uint8)
    start, end = np.frombuffer(raw[:16].tobytes(), np.uint64)
    L, n = np.frombuffer(raw[16:24].tobytes(), np.uint32)
    assert (start, L, n) == (0, 32, sa._idx.n) and end == n
    assert np.array_equal(np.frombuffer(raw[24:].tobytes(), np.uint32), sa.suffix_array)
    sa.close()
    sb.close()
    # CSV index: file offsets and quoted bits survive the round trip
    data = W.csv_file_bytes(400, seed=5, quoted_fraction=0.3)
    path = str(tmp_path / "t.csv")
    with open(path, "wb") as f:
        f.write(data)
    a = SuffixArray(csv_file=path, search_column="name", max_suffix_length=32, csv_query="text")
    a.save(str(tmp_path / "csv.sa"))
    b = SuffixArray.load(str(tmp_path / "csv.sa"))
    assert np.array_equal(a.suffix_array, b.suffix_array)
    assert np.array_equal(a._idx.quoted_bits(), b._idx.quoted_bits())
    assert a.query_records("inc", k=20) == b.query_records("inc", k=20)
    with pytest.raises(Exception):
        S.Index(16).load(str(tmp_path / "csv.sa"))        # max_suffix_length mismatch
    a.close()
    b.close()


def test_two_devices_in_one_process():
    """launch configuration (dynamic shared-memory opt-in, occupancy grid) is cached per
    DEVICE, so a second index on another GPU of the same process builds too."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs in one process")
    rng = np.random.default_rng(5)
    t = (rng.integers(0, 4, size=300_000) + ord("a")).astype(np.uint8)
    t[rng.random(t.size) < 0.02] = 10
    want = O.oracle_build(t, 32)
    for dev in (0, 1, 0):
        idx = S.Index(32, device=dev)
        idx.build_text(t)
        assert np.array_equal(idx.sa(), want), dev
        idx.close()


def test_load_rejects_truncated_and_corrupt_containers(tmp_path):
    """sa_load checks the file length against the header and that every stored position is in
    range (a corrupt container must fail with an error, not become an out-of-bounds device read)."""
    rng = np.random.default_rng(3)
    t = (rng.integers(0, 6, size=50_000) + ord("a")).astype(np.uint8)
    t[rng.random(t.size) < 0.03] = 10
    idx = S.Index(32)
    idx.build_text(t)
    p = str(tmp_path / "x.sa")
    idx.save(p)
    want = idx.sa()
    idx.close()
    raw = bytearray(open(p, "rb").read())
    ok = S.Index(32)
    ok.load(p)
    assert np.array_equal(ok.sa(), want)
    ok.close()
    with open(str(tmp_path / "short.sa"), "wb") as f:
        f.write(raw[: len(raw) // 2])
    with pytest.raises(S.SaCudaError):
        S.Index(32).load(str(tmp_path / "short.sa"))
    bad = bytearray(raw)
    bad[-4:] = (0xFFFFFFF0).to_bytes(4, "little")          # the last array entry now points far outside the text
    with open(str(tmp_path / "bad.sa"), "wb") as f:
        f.write(bad)
    with pytest.raises(S.SaCudaError):
        S.Index(32).load(str(tmp_path / "bad.sa"))
