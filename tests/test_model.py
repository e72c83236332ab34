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
