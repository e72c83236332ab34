"""CSV files of 2 GiB and more: one suffix array per 2 GiB partition of the FILE, each partition resuming
where the previous one stopped (engine.c:1437-1438, :1469-1481) -- sa_build_csv_partition.

The always-on test checks the partition entry against sa_build_csv / the oracle on a small file; the
full-size test (a 2.3 GB file generated on the box, the reference's two partitions built by the
unmodified engine.c, arrays and quoted bits compared bit for bit per partition) needs about four minutes
of CPU and 20 GB of host memory: SA_TEST_BIG_CSV=1 enables it (scripts/gpu_big_csv.sh runs it)."""
import os
import tempfile

import numpy as np
import pytest

from oracle import pyoracle as O
from suffixarray_b200 import _lib as S
from suffixarray_b200 import workloads as W

pytestmark = pytest.mark.gpu


def test_partition_entry_on_a_small_file(tmp_path):
    data = W.csv_file_bytes(3000, seed=4, quoted_fraction=0.2)
    p = str(tmp_path / "s.csv")
    with open(p, "wb") as f:
        f.write(data)
    a, b = S.Index(32), S.Index(32)
    a.build_csv(p, "name")
    end, size = b.build_csv_partition(p, "name", 0)
    assert size == len(data) and end >= size
    assert np.array_equal(a.sa(), b.sa()) and np.array_equal(a.quoted_bits(), b.quoted_bits())
    assert np.array_equal(a.text(), b.text())
    qs = [b"inc", b"llc", b"a", b"corp"]
    ra, rb = a.result(), b.result()
    for flags in (0, S.SA_QUERY_FILE_SEMANTICS):
        x, y = ra.ranges(qs, 10, flags), rb.ranges(qs, 10, flags)
        assert all(np.array_equal(u, v) for u, v in zip(x, y))
        assert all(np.array_equal(u, v) for u, v in zip(ra.records(), rb.records()))
    # a start in the middle of the file: the scan begins there with column 0, like a second partition would
    mid = len(data) // 2 + 7
    c = S.Index(32)
    end2, _ = c.build_csv_partition(p, "name", mid)
    buf = np.zeros(len(data) + 8, np.uint8)
    buf[: len(data)] = np.frombuffer(data, np.uint8)
    text, mp, qd = O.oracle_csv_scan(data[mid:], 1, 4)
    assert c.n == text.size and np.array_equal(c.text(), text)
    sa = O.oracle_build(text, 32)
    sa_file, q_rank = O.oracle_csv_remap(sa, mp + np.uint32(mid), qd)
    assert np.array_equal(c.sa(), sa_file) and np.array_equal(c.quoted_bits(), q_rank)
    for ix in (a, b, c):
        ix.close()


@pytest.mark.skipif(os.environ.get("SA_TEST_BIG_CSV") != "1", reason="set SA_TEST_BIG_CSV=1 (2.3 GB file, ~5 min of CPU)")
def test_two_partitions_match_the_reference():
    if not O.have_ref():
        pytest.skip("oracle/_ref/libsa_ref.so not present")
    d = tempfile.mkdtemp(prefix="sa_big_")
    path = os.path.join(d, "big.csv")
    try:
        rows, fbytes = W.write_csv_file(path, 1500, seed=7001, quoted_rows_per_chunk=5000)
        assert (1 << 31) < fbytes < (1 << 32) - 4096
        O.ref_use_all_cores()
        R = O.RefCsvIndex(path, "name", 32)
        nparts = R.partitions()
        assert nparts == 2
        from suffixarray_b200 import SuffixArray
        sa = SuffixArray(csv_file=path, search_column="name", max_suffix_length=32)
        arrays = sa.suffix_array
        assert len(arrays) == nparts
        start = 0
        for p in range(nparts):
            rs, re_, rsa, rq = R.partition(p)
            assert rs == start
            assert arrays[p].size == rsa.size, (p, arrays[p].size, rsa.size)
            assert np.array_equal(arrays[p], rsa), p
            assert np.array_equal(sa._parts[p].quoted_bits(), rq), p
            start = re_
        # queries walk the partitions; every returned row really holds the query
        for q in ("inc", "holdings", "zzzzqq"):
            recs = sa.query_records(q, k=20)
            assert all(q in r["name"].lower() for r in recs)
        sa.close()
        R.close()
    finally:
        try:
            os.unlink(path)
            os.rmdir(d)
        except OSError:
            pass
