"""GPU parity, build: libsa_cuda.so (through the C ABI) vs the oracle, bit for bit.

Mirrors what a test of construct_truncated_suffix_array (engine.c:837-866) would assert if the
reference had one: the produced uint32 array equals ORDER-E on the same input.
"""
import numpy as np
import pytest

from conftest import GRID_L, GRID_N, GRID_SIGMA, random_text
from oracle import pyoracle as O
from suffixarray_b200 import _lib as S
from suffixarray_b200 import workloads as W

pytestmark = pytest.mark.gpu


def gpu_build(text, L, lowercase=False, **dbg):
    idx = S.Index(L)
    for k, v in dbg.items():
        idx.debug_set(k, v)
    idx.build_text(text, lowercase=lowercase)
    sa = idx.sa()
    st = idx.stats()
    idx.close()
    return sa, st


def assert_same(text, L, **dbg):
    want = O.oracle_build(text, L)
    got, st = gpu_build(text, L, **dbg)
    if not np.array_equal(want, got):
        bad = np.flatnonzero(want != got)
        raise AssertionError(f"n={text.size} L={L} dbg={dbg}: {bad.size} ranks differ, first at {bad[:5]} "
                             f"want {want[bad[:5]]} got {got[bad[:5]]} stats={st}")
    return st


def test_banana_known_answer():
    # SURVEY 3.2 known-answer vector (produced by the reference itself)
    sa, _ = gpu_build(b"banana\nbandana", 32)
    assert sa.tolist() == [6, 13, 5, 11, 3, 1, 8, 0, 7, 10, 12, 4, 2, 9]


def test_empty_text():
    idx = S.Index(32)
    idx.build_text(b"")
    assert idx.n == 0 and idx.sa().size == 0
    idx.close()


@pytest.mark.parametrize("sigma", GRID_SIGMA)
def test_grid(sigma):
    rng = np.random.default_rng(100 + sigma)
    for n in GRID_N:
        for L in GRID_L:
            assert_same(random_text(rng, n, sigma), L)


@pytest.mark.parametrize("L", [1, 7, 8, 9, 16, 17, 32, 64, 100])
def test_edge_texts(L):
    rng = np.random.default_rng(7)
    cases = [
        np.full(5000, ord("a"), np.uint8),                       # one symbol: ties through every round
        np.full(777, 10, np.uint8),                              # only newlines (never refined)
        np.full(500, 255, np.uint8),                             # only 0xFF (never refined)
        np.tile(np.frombuffer(b"abcabcabd\n", np.uint8), 700),   # long repeats
        np.tile(np.frombuffer(b"x" * 70 + b"\n", np.uint8), 90), # common prefixes longer than L
        rng.integers(0, 256, 9000).astype(np.uint8),             # all byte values incl. NUL / 0xFF
        rng.integers(0, 3, 6000).astype(np.uint8) * 127,         # {0, 127, 254}
        np.frombuffer("naïve café\nüber straße\n日本語テキスト\n".encode() * 50, np.uint8),  # UTF-8
        np.frombuffer(b"\n\n\nab\n\n\nab\n\n" * 40, np.uint8),   # empty documents
        np.concatenate([np.full(40, ord("q"), np.uint8), rng.integers(97, 99, 3000).astype(np.uint8)]),
    ]
    for t in cases:
        assert_same(t, L)


def test_0xff_and_newline_buckets():
    rng = np.random.default_rng(11)
    t = rng.choice(np.array([10, 255, 97, 98], np.uint8), size=20000, p=[0.2, 0.2, 0.3, 0.3])
    for L in (4, 16, 32):
        assert_same(t, L)


@pytest.mark.parametrize("gbits", [0, 40, 50, 56])
def test_small_refinement_words(gbits):
    # force wide group ids so refinement rounds extend by 1-3 bytes only
    rng = np.random.default_rng(13)
    t = random_text(rng, 30000, 2, newline_p=0.01)
    assert_same(t, 32, force_gbits=gbits)


def test_lowercase_flag():
    text, _ = W.company_names_joined(3000, seed=5, upper=True)
    want = O.oracle_build(O.oracle_lowercase(text), 32)
    got, _ = gpu_build(text, 32, lowercase=True)
    assert np.array_equal(want, got)


@pytest.mark.parametrize("L", [8, 16, 32, 64])
def test_company_names_200k(L):
    text, _ = W.company_names_joined(200_000, seed=21, upper=False)
    st = assert_same(text, L)
    assert st["finish_runs"] > 0


@pytest.mark.parametrize("L", [8, 32, 64])
def test_english_4m(L):
    text = W.english_documents(4_000_000, seed=22)
    assert_same(text, L)


def test_rebuild_reuses_index():
    rng = np.random.default_rng(17)
    idx = S.Index(32)
    for n in (50000, 100, 80000, 0, 33):
        t = random_text(rng, n, 4)
        idx.build_text(t)
        assert np.array_equal(idx.sa(), O.oracle_build(t, 32))
    idx.close()


def test_i64_export_and_text_roundtrip():
    rng = np.random.default_rng(19)
    t = random_text(rng, 10000, 8)
    idx = S.Index(32)
    idx.build_text(t)
    assert np.array_equal(idx.sa_i64(), idx.sa().astype(np.int64))
    assert np.array_equal(idx.text(), t)
    idx.close()


def test_large_properties_20m():
    """Full-size style check without the oracle: permutation + key order + spot re-derivation."""
    text, _ = W.company_names_joined(700_000, seed=23, upper=False)
    n = text.size
    sa, st = gpu_build(text, 32)
    assert sa.size == n
    seen = np.zeros(n, dtype=bool)
    seen[sa] = True
    assert seen.all()
    # first byte is non-decreasing (depth-0 counting sort, engine.c:734-758)
    fb = text[sa]
    assert (np.diff(fb.astype(np.int16)) >= 0).all()
    # and the whole array equals the oracle (20 MB finishes in a few seconds)
    assert np.array_equal(sa, O.oracle_build(text, 32))


@pytest.mark.parametrize("code_bits", [8, 7, 5, 3])
def test_alphabet_code_widths(code_bits):
    """The compressed-alphabet keys (code_bits < 8) and the raw-byte keys (8) give the same SA."""
    rng = np.random.default_rng(29)
    cases = [
        (random_text(rng, 30000, 2, newline_p=0.02), 32),
        (random_text(rng, 30000, 3), 17),
        (random_text(rng, 5000, 1, newline_p=0.01), 64),
        (W.english_documents(300_000, seed=31), 32),
        (W.company_names_joined(20_000, seed=33, upper=False)[0], 32),
        (np.frombuffer(b"\n\n\nab\n\n\nab\n\n" * 400, np.uint8), 16),
        (rng.choice(np.array([10, 255, 0, 97, 98], np.uint8), size=20000), 32),
    ]
    for t, L in cases:
        st = assert_same(t, L, code_bits=code_bits)
        assert st["code_bits"] >= min(code_bits, 8) or code_bits == 8


def test_alphabet_auto_widths():
    rng = np.random.default_rng(37)
    for sigma, want_bits in ((1, 2), (2, 2), (6, 3), (26, 5), (100, 7), (200, 8)):
        t = random_text(rng, 20000, sigma, base=20 if sigma > 26 else ord("a"))
        st = assert_same(t, 32)
        assert st["code_bits"] == want_bits, (sigma, st["code_bits"])
