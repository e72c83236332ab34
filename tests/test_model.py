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
