"""The NumPy model of the GPU's flat-pass decomposition (oracle/model.py) equals the recursive
restatement (oracle/sa_oracle.c) -- the argument that pack -> LSD sort -> tie refinement ->
<=32-bucket finish reproduces recursive_bucket_sort (engine.c:696-812)."""
import numpy as np
import pytest

from conftest import random_text
from oracle import model as M
from oracle import pyoracle as O


@pytest.mark.parametrize("sigma", [1, 2, 3, 8, 26])
def test_model_equals_oracle(sigma):
    rng = np.random.default_rng(200 + sigma)
    for n in (1, 2, 31, 32, 33, 34, 200, 3000):
        for L in (1, 4, 8, 15, 16, 17, 31, 32, 33, 64):
            t = random_text(rng, n, sigma)
            assert np.array_equal(M.model_build(t, L), O.oracle_build(t, L)), (n, sigma, L)


def test_model_terminators_and_binary():
    rng = np.random.default_rng(9)
    t = rng.choice(np.array([10, 255, 0, 97, 98], np.uint8), size=6000)
    for L in (4, 16, 32):
        assert np.array_equal(M.model_build(t, L), O.oracle_build(t, L))


def test_csv_scan_model_equals_oracle_or_flags_the_file():
    """The device column scanner's algorithm (oracle/model.py csv_scan_model <-> csrc/csv_scan.cu):
    files written by Python's csv module are regular and give the oracle's streams; on files with
    stray quotes the model either raises an irregularity flag (the library then takes the
    reference-shaped host scan) or still agrees with the oracle -- never a silent difference."""
    import csv
    import io
    from oracle import model as M
    rng = np.random.default_rng(7)
    alpha = list('abXY ,"\n.')
    regular = irregular = 0
    for trial in range(400):
        ncols = int(rng.integers(1, 5))
        if trial % 2 == 0:                                   # RFC-4180 file
            rows = [[f"c{i}" for i in range(ncols)]] + [
                ["".join(rng.choice(alpha, size=int(rng.integers(0, 8)))) for _ in range(ncols)]
                for _ in range(int(rng.integers(1, 25)))]
            buf = io.StringIO()
            csv.writer(buf, lineterminator="\n").writerows(rows)
            data = buf.getvalue().encode()
            if trial % 6 == 0:
                data = data[:-1]
        else:                                                # raw bytes: quotes anywhere
            data = "".join(rng.choice(alpha, size=int(rng.integers(1, 120)))).encode()
        for col in range(ncols):
            flags, text, mp = M.csv_scan_model(data, col, ncols)
            if trial % 2 == 0:
                assert flags == 0, (trial, data)
            if flags:
                irregular += 1
                continue
            regular += 1
            want_text, want_map, want_q = O.oracle_csv_scan(data, col, ncols)
            assert np.array_equal(text, want_text), (trial, col, data)
            assert np.array_equal(mp & 0x7FFFFFFF, want_map), (trial, col, data)
            assert np.array_equal((mp >> 31).astype(np.uint8), want_q), (trial, col, data)
    assert regular > 300 and irregular > 50


def test_finish_per_rank_rules_and_128_record_halos():
    """(a) finish_kernel's per-rank bucket rule (head iff lcp < d*) gives the sequential walk's result;
    (b) the sharded finish: a shard extended by min(128, what exists) records on either side, finished
    on its own, reproduces its slice of the global result -- for random shard boundaries, including
    ones inside tied buckets and shards shorter than the halo (sa_shard.cu, sharded.EDGE)."""
    rng = np.random.default_rng(21)
    EDGE = 128
    for sigma, n, L in ((2, 3000, 32), (3, 5000, 16), (4, 4000, 64), (26, 6000, 32), (2, 700, 8), (3, 90, 32)):
        text = random_text(rng, n, sigma, newline_p=0.03)
        lprime = min(L, n)
        T = M.padded(text)
        SA, lcp = M.refine_state(T, n, lprime)
        want = M.finish(T, SA, lcp, n, lprime)
        assert np.array_equal(M.finish_flat(T, SA, lcp, n, lprime), want), (sigma, n, L)
        assert np.array_equal(want.astype(np.uint32), O.oracle_build(text, L))
        cuts = sorted(set([0, n] + rng.integers(0, n + 1, size=12).tolist() + [n // 2, n // 2 + 1, n // 2 + 40]))
        for a, b in zip(cuts[:-1], cuts[1:]):
            hl, hr = min(EDGE, a), min(EDGE, n - b)
            ext_sa, ext_lcp = SA[a - hl: b + hr].copy(), lcp[a - hl: b + hr].copy()
            got = M.finish_flat(T, ext_sa, ext_lcp, ext_sa.size, lprime, lo=hl, hi=hl + (b - a))
            assert np.array_equal(got[hl: hl + (b - a)], want[a:b]), (sigma, n, L, a, b)
