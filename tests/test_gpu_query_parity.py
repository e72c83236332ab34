"""GPU parity, query: sa_query_ranges / sa_query_hits vs the oracle (and vs the reference's own
get_substring_positions where oracle/_ref is present), bit for bit: same inclusive rank range with
the UINT32_MAX sentinels (engine.c:869-918), same first min(k, count) hits in rank order with
duplicates kept (engine.c:1185-1187), record index = number of '\\n' before the hit."""
import numpy as np
import pytest

from conftest import random_text, safe_queries_for
from oracle import pyoracle as O
from suffixarray_b200 import _lib as S
from suffixarray_b200 import workloads as W

pytestmark = pytest.mark.gpu


def check_queries(text, L, queries, k=1000, ref=False):
    idx = S.Index(L)
    idx.build_text(text)
    sa = idx.sa()
    want_sa = O.oracle_build(text, L)
    assert np.array_equal(sa, want_sa)
    ws, we = O.oracle_query(text, want_sa, L, queries)
    # one thread per query (with the probe cache of the top levels off / shallow / automatic) and eight lanes per query
    for mode, levels in ((1, 0), (1, 2), (1, 7), (1, -1), (2, -1)):
        idx.debug_set("query_kernel", mode)
        idx.debug_set("qtop_levels", levels)
        s, e, off = idx.query_ranges(queries, k=k)
        assert np.array_equal(s, ws), f"kernel {mode}/{levels}: start differs at {np.flatnonzero(s != ws)[:5]}"
        assert np.array_equal(e, we), f"kernel {mode}/{levels}: end differs at {np.flatnonzero(e != we)[:5]}"
    woff, wpos, wdoc = O.oracle_hits(text, want_sa, ws, we, k)
    assert np.array_equal(off, woff)
    pos, doc = idx.query_hits(int(off[-1]))
    assert np.array_equal(pos, wpos)
    assert np.array_equal(doc, wdoc)
    if ref and O.have_ref():
        rs, re_, skipped = O.ref_query(text, want_sa, L, queries)
        keep = skipped == 0
        assert keep.sum() > 0
        assert np.array_equal(s[keep], rs[keep]) and np.array_equal(e[keep], re_[keep])
    idx.close()
    return s, e, off


def test_banana_known_answer():
    # get_substring_positions("ana") = {3,5}, ("ban") = {7,8} on the reference
    s, e, off = check_queries(np.frombuffer(b"banana\nbandana", np.uint8), 32, [b"ana", b"ban", b"zzz", b"b"])
    assert (s[0], e[0]) == (3, 5) and (s[1], e[1]) == (7, 8)
    assert s[2] == 0xFFFFFFFF and e[2] == 0xFFFFFFFF


@pytest.mark.parametrize("sigma,L", [(2, 8), (3, 16), (8, 32), (26, 32), (26, 64), (4, 4)])
def test_random_texts(sigma, L):
    rng = np.random.default_rng(1000 + sigma * 7 + L)
    for n in (40, 500, 20000):
        t = random_text(rng, n, sigma)
        t[0] = 10                     # rank 0 is a newline suffix: the reference never wraps
        qs = safe_queries_for(t, rng, 300, max_len=70)
        check_queries(t, L, qs, k=50, ref=True)


def test_company_names_whole_name_queries():
    text, starts = W.company_names_joined(100_000, seed=31, upper=False)
    qs = W.whole_name_queries(text, starts, 3000, seed=32)
    s, e, off = check_queries(text, 32, qs, k=1000, ref=True)
    assert (s != 0xFFFFFFFF).all()          # every sampled name occurs


def test_english_substring_queries():
    text = W.english_documents(3_000_000, seed=33)
    qb, qo = W.substring_queries(text, 20000, seed=34)
    qs = W.unpack_queries(qb, qo)
    check_queries(text, 32, qs, k=1000, ref=True)


def test_query_edge_cases():
    text = np.frombuffer(b"\nalpha beta\ngamma\n\ndelta alpha\nalp", np.uint8)
    qs = [b"alpha", b"alp", b"a", b"zz", b"delta alpha", b"x" * 100, b"alpha\0beta", b"\0", b"", b"gamma\n"]
    check_queries(text, 32, qs, k=2)
    check_queries(text, 3, qs, k=1000)       # queries longer than L are truncated to L (engine.c:881)
    check_queries(text, 32, qs, k=0)


def test_k_limits_hits_and_keeps_duplicates():
    text = np.frombuffer(b"\n" + b"abab abab\n" * 50, np.uint8)
    s, e, off = check_queries(text, 32, [b"ab", b"abab", b"b a"], k=7)
    assert off.tolist() == [0, 7, 14, 21]


def test_empty_index_and_empty_batch():
    idx = S.Index(32)
    idx.build_text(b"")
    s, e, off = idx.query_ranges([b"abc"], k=10)
    assert s[0] == 0xFFFFFFFF and e[0] == 0xFFFFFFFF and off.tolist() == [0, 0]
    idx.build_text(b"\nabc\nabd")
    s, e, off = idx.query_ranges([], k=10)
    assert s.size == 0 and off.tolist() == [0]
    idx.close()


def test_doc_starts():
    text = np.frombuffer(b"ab\ncd\n\nef", np.uint8)
    idx = S.Index(32)
    idx.build_text(text)
    assert idx.num_docs == 4
    assert idx.doc_starts().tolist() == [0, 3, 6, 7]
    idx.close()
