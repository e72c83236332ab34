"""Golden vectors produced by RUNNING the reference (tests/golden/make_golden.py, engine.c compiled
unmodified).  CPU: the oracle port must reproduce them.  GPU: libsa_cuda.so must reproduce them."""
import glob
import os
import tempfile

import numpy as np
import pytest

from oracle import pyoracle as O
from suffixarray_b200 import workloads as W

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TEXT_CASES = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLD, "*.npz")) if "csv" not in p)


def load(name):
    return np.load(os.path.join(GOLD, name + ".npz"))


def test_golden_files_present():
    assert len(TEXT_CASES) >= 10 and os.path.exists(os.path.join(GOLD, "csv_L32.npz"))


@pytest.mark.parametrize("name", TEXT_CASES)
def test_oracle_matches_golden(name):
    g = load(name)
    L = int(g["L"])
    sa = O.oracle_build(g["text"], L)
    assert np.array_equal(sa, g["sa"])
    qs = W.unpack_queries(g["qbytes"], g["qoff"])
    s, e = O.oracle_query(g["text"], sa, L, qs)
    keep = g["skipped"] == 0
    assert np.array_equal(s[keep], g["start"][keep]) and np.array_equal(e[keep], g["end"][keep])


def test_oracle_csv_matches_golden():
    g = load("csv_L32")
    data = g["csv_bytes"].tobytes()
    for col, name in ((1, "name"), (2, "city")):
        text, mp, qd = O.oracle_csv_scan(data, col, 4)
        sa = O.oracle_build(text, 32)
        sa_file, q_rank = O.oracle_csv_remap(sa, mp, qd)
        assert np.array_equal(sa_file, g["sa_" + name])
        assert np.array_equal(q_rank, g["quoted_" + name])


@pytest.mark.gpu
@pytest.mark.parametrize("name", TEXT_CASES)
def test_gpu_matches_golden(name):
    from suffixarray_b200 import _lib as S
    g = load(name)
    L = int(g["L"])
    idx = S.Index(L)
    idx.build_text(g["text"])
    assert np.array_equal(idx.sa(), g["sa"])
    s, e, _ = idx.query_ranges_packed(g["qbytes"], g["qoff"], k=1000)
    keep = g["skipped"] == 0
    assert np.array_equal(s[keep], g["start"][keep]) and np.array_equal(e[keep], g["end"][keep])
    idx.close()


@pytest.mark.gpu
def test_gpu_csv_matches_golden():
    from suffixarray_b200 import _lib as S
    g = load("csv_L32")
    with tempfile.NamedTemporaryFile(suffix=".csv", delete=False) as f:
        f.write(g["csv_bytes"].tobytes())
        path = f.name
    try:
        for name in ("name", "city"):
            idx = S.Index(32)
            idx.build_csv(path, name)
            assert np.array_equal(idx.sa(), g["sa_" + name])
            assert np.array_equal(idx.quoted_bits(), g["quoted_" + name])
            idx.close()
    finally:
        os.unlink(path)
