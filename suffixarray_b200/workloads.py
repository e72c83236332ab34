"""Seeded synthetic workloads of the shapes BASELINE.json names.

Pure NumPy (host side only; nothing here is on the timed path).  The reference ships no data
(its tests read data/companies-2023-q4-sm.csv, which is not in the repository,
tests/test.py:235), so every shape is generated:

  C1  company-name strings, upper-cased as tests/test.py:63 does, whole-name queries
      (tests/test.py:65-66)
  C2  English-like document list + substring queries (80 % present, 20 % one byte mutated)
  C3  RFC-4180 CSV `id,name,city,score` with a share of quoted `name` cells
"""
from __future__ import annotations

import numpy as np

_SYLLABLES = [
    "ka", "lo", "mi", "ra", "ten", "vor", "sa", "ne", "qui", "dor", "al", "be", "cor", "du", "el", "fa",
    "gri", "ha", "in", "jo", "kel", "lu", "mar", "no", "or", "pa", "ri", "sol", "tra", "ul", "ven", "wa",
    "xe", "yo", "za", "bri", "cle", "dra", "eth", "flo", "gan", "hel", "ix", "jan", "kor", "lin", "mon", "nor",
    "op", "pre", "qua", "ros", "ste", "tur", "um", "vi", "wes", "xi", "yar", "zen", "ab", "ec", "id", "ox",
]
_SUFFIXES = ["inc", "llc", "ltd", "corp", "group", "holdings", "co", "gmbh", "sa", "plc"]


def _zipf_indices(rng: np.random.Generator, size: int, vocab: int, a: float) -> np.ndarray:
    """Zipf(a) over ranks 1..vocab by inverse-CDF lookup (bounded, unlike rng.zipf)."""
    w = 1.0 / np.power(np.arange(1, vocab + 1, dtype=np.float64), a)
    cdf = np.cumsum(w)
    cdf /= cdf[-1]
    return np.searchsorted(cdf, rng.random(size), side="left").astype(np.int64)


def _flatten_vocab(words: list[bytes]):
    lens = np.fromiter((len(w) for w in words), dtype=np.int64, count=len(words))
    off = np.zeros(len(words) + 1, dtype=np.int64)
    np.cumsum(lens, out=off[1:])
    flat = np.frombuffer(b"".join(words), dtype=np.uint8)
    return flat, off, lens


def _assemble(flat, voff, vlen, tok_word, sep_bytes, sep_len):
    """Concatenate tokens word[tok_word[i]] + sep_bytes[i, :sep_len[i]] into one u8 array."""
    wl = vlen[tok_word]
    tl = wl + sep_len
    start = np.zeros(tok_word.size + 1, dtype=np.int64)
    np.cumsum(tl, out=start[1:])
    total = int(start[-1])
    tok_of = np.repeat(np.arange(tok_word.size, dtype=np.int64), tl)
    within = np.arange(total, dtype=np.int64) - start[:-1][tok_of]
    wl_of = wl[tok_of]
    in_word = within < wl_of
    src = voff[tok_word][tok_of] + np.minimum(within, wl_of - 1)
    out = flat[src]
    sidx = np.clip(within - wl_of, 0, sep_bytes.shape[1] - 1)
    out = np.where(in_word, out, sep_bytes[tok_of, sidx])
    return out.astype(np.uint8), start


# ------------------------------------------------------------------------------------- C1
def company_names(num: int, seed: int = 1001, mean_tokens: float = 2.5, upper: bool = True) -> list[bytes]:
    """`num` company-like names: 1-4 Zipf(1.2) tokens from a 30 000-word syllable vocabulary,
    35 % with a legal-form suffix; upper-cased (tests/test.py:63) so lower-casing is exercised."""
    text, starts = company_names_joined(num, seed, upper)
    b = text.tobytes()
    return [b[starts[i]:starts[i + 1] - 1] for i in range(num)]


def company_names_joined(num: int, seed: int = 1001, upper: bool = True):
    """-> (u8 text = names joined by '\\n' WITHOUT a trailing '\\n', i64 starts[num+1]) where
    name i is text[starts[i] : starts[i+1]-1]."""
    rng = np.random.default_rng(np.random.PCG64(seed))
    nv = 30000
    syl = [s.encode() for s in _SYLLABLES]
    k = rng.integers(2, 6, size=nv)
    picks = rng.integers(0, len(syl), size=(nv, 5))
    vocab = [b"".join(syl[picks[i, j]] for j in range(k[i])) for i in range(nv)]
    vocab += [s.encode() for s in _SUFFIXES]
    flat, voff, vlen = _flatten_vocab(vocab)
    ntok = rng.integers(1, 5, size=num)
    has_suffix = rng.random(num) < 0.35
    per = ntok + has_suffix
    total_tok = int(per.sum())
    tok_word = _zipf_indices(rng, total_tok, nv, 1.2)
    last = np.cumsum(per) - 1
    tok_word[last[has_suffix]] = nv + rng.integers(0, len(_SUFFIXES), size=int(has_suffix.sum()))
    sep = np.full((total_tok, 1), ord(" "), dtype=np.uint8)
    sep[last, 0] = ord("\n")
    sep_len = np.ones(total_tok, dtype=np.int64)
    text, tstart = _assemble(flat, voff, vlen, tok_word, sep, sep_len)
    if upper:
        lo = (text >= ord("a")) & (text <= ord("z"))
        text = np.where(lo, text - 32, text).astype(np.uint8)
    name_start = np.zeros(num + 1, dtype=np.int64)
    name_start[1:] = tstart[1:][last]
    return text[:-1].copy(), name_start


def whole_name_queries(text: np.ndarray, starts: np.ndarray, nq: int, seed: int = 1002) -> list[bytes]:
    """nq whole names sampled without replacement (tests/test.py:65-66), lower-cased as
    query_records does (suffix_array.pyx:228)."""
    rng = np.random.default_rng(np.random.PCG64(seed))
    num = starts.size - 1
    pick = rng.choice(num, size=min(nq, num), replace=False)
    b = text.tobytes()
    return [b[starts[i]:starts[i + 1] - 1].lower() for i in pick]


# ------------------------------------------------------------------------------------- C2
def english_documents(n_chars: int, seed: int = 2001) -> np.ndarray:
    """English-like text of ~n_chars bytes (exactly n_chars after trimming): 50 000-word
    vocabulary (length Poisson(4)+1 over a-z), Zipf(1.07) unigram draw, sentences of 5-25 words
    joined by '. ', documents of 200-2000 chars joined by '\\n' (no trailing '\\n')."""
    rng = np.random.default_rng(np.random.PCG64(seed))
    nv = 50000
    wl = rng.poisson(4.0, size=nv) + 1
    letters = rng.integers(ord("a"), ord("z") + 1, size=int(wl.sum()), dtype=np.int64).astype(np.uint8)
    voff = np.zeros(nv + 1, dtype=np.int64)
    np.cumsum(wl, out=voff[1:])
    mean_tok = float((wl * (1.0 / np.arange(1, nv + 1) ** 1.07)).sum() / (1.0 / np.arange(1, nv + 1) ** 1.07).sum()) + 1.1
    out_parts = []
    have = 0
    while have < n_chars + 1:
        need = n_chars + 1 - have
        ntok = int(need / mean_tok * 1.05) + 1024
        tok_word = _zipf_indices(rng, ntok, nv, 1.07)
        # sentence ends: geometric-ish with lengths uniform 5..25 -> mark by cumulative draw
        sent_len = rng.integers(5, 26, size=ntok // 5 + 2)
        sent_end = np.cumsum(sent_len) - 1
        sent_end = sent_end[sent_end < ntok]
        sep = np.full((ntok, 2), ord(" "), dtype=np.uint8)
        sep_len = np.ones(ntok, dtype=np.int64)
        sep[sent_end, 0] = ord(".")
        sep_len[sent_end] = 2
        # document ends: at the sentence end nearest to a running 200..2000 char budget
        tl = wl[tok_word] + sep_len
        cum = np.cumsum(tl)
        doc_len = rng.integers(200, 2001, size=int(cum[-1] // 200) + 2)
        doc_cut = np.cumsum(doc_len)
        doc_cut = doc_cut[doc_cut < cum[-1]]
        se_cum = cum[sent_end]
        j = np.searchsorted(se_cum, doc_cut, side="left")
        j = np.unique(j[j < sent_end.size])
        de = sent_end[j]
        sep[de, 1] = ord("\n")          # ".\n"
        part, _ = _assemble(letters, voff, wl, tok_word, sep, sep_len)
        out_parts.append(part)
        have += part.size
    text = np.concatenate(out_parts)[:n_chars].copy()
    if text.size and text[-1] == ord("\n"):
        text[-1] = ord(".")
    return text


def substring_queries(text: np.ndarray, nq: int, seed: int = 2002, min_len: int = 3, max_len: int = 64,
                      miss_fraction: float = 0.2):
    """-> (qbytes u8, qoff u64[nq+1]).  80 % are text substrings at uniform positions with
    uniform length in [min_len, max_len], re-drawn while they contain '\\n'; the remaining
    share has one byte replaced by a random letter (mostly misses).  Text must be lower-case
    already (queries are lower-cased by query_records, suffix_array.pyx:228)."""
    rng = np.random.default_rng(np.random.PCG64(seed))
    n = text.size
    lens = rng.integers(min_len, max_len + 1, size=nq).astype(np.int64)
    lens = np.minimum(lens, max(1, n - 1))
    pos = rng.integers(0, np.maximum(1, n - lens), size=nq).astype(np.int64)
    nl = np.flatnonzero(text == ord("\n")).astype(np.int64)
    for _ in range(64):
        nxt = np.searchsorted(nl, pos, side="left")
        bad = (nxt < nl.size) & (nl[np.minimum(nxt, max(nl.size - 1, 0))] < pos + lens) if nl.size else np.zeros(nq, bool)
        nb = int(bad.sum())
        if nb == 0:
            break
        pos[bad] = rng.integers(0, np.maximum(1, n - lens[bad]), size=nb)
    else:
        lens[bad] = 1      # pathological text (newlines everywhere): fall back to 1-byte queries
        pos[bad] = np.flatnonzero(text != ord("\n"))[:1].repeat(nb) if (text != ord("\n")).any() else 0
    qoff = np.zeros(nq + 1, dtype=np.uint64)
    np.cumsum(lens, out=qoff[1:].view(np.int64))
    total = int(qoff[-1])
    q_of = np.repeat(np.arange(nq, dtype=np.int64), lens)
    within = np.arange(total, dtype=np.int64) - qoff[:-1].astype(np.int64)[q_of]
    qbytes = text[pos[q_of] + within].copy()
    mut = np.flatnonzero(rng.random(nq) < miss_fraction)
    if mut.size:
        at = qoff[:-1].astype(np.int64)[mut] + rng.integers(0, lens[mut])
        qbytes[at] = rng.integers(ord("a"), ord("z") + 1, size=mut.size).astype(np.uint8)
    return qbytes, qoff


def fixed_length_queries(text: np.ndarray, nq: int, m: int, seed: int):
    """C5 sweep: nq text substrings of exactly m bytes (seed 5000+m in the sweep)."""
    return substring_queries(text, nq, seed=seed, min_len=m, max_len=m, miss_fraction=0.0)


def unpack_queries(qbytes: np.ndarray, qoff: np.ndarray) -> list[bytes]:
    b = qbytes.tobytes()
    o = qoff.astype(np.int64)
    return [b[o[i]:o[i + 1]] for i in range(o.size - 1)]


# ------------------------------------------------------------------------------------- C3
def csv_file_bytes(num_rows: int, seed: int = 3001, quoted_fraction: float = 0.05) -> bytes:
    """RFC-4180 CSV `id,name,city,score`; `quoted_fraction` of the name cells are quoted and
    contain an embedded ',' and a doubled '""'."""
    rng = np.random.default_rng(np.random.PCG64(seed))
    names = company_names(num_rows, seed=seed + 1, upper=False)
    cities = [b"springfield", b"riverton", b"lakeside", b"fairview", b"georgetown", b"salem", b"madison"]
    quoted = rng.random(num_rows) < quoted_fraction
    city = rng.integers(0, len(cities), size=num_rows)
    score = rng.integers(0, 1000, size=num_rows)
    rows = [b"id,name,city,score"]
    for i in range(num_rows):
        nm = names[i]
        if quoted[i]:
            half = len(nm) // 2
            nm = b'"' + nm[:half] + b', ""' + nm[half:] + b'"" co"'
        rows.append(b"%d,%s,%s,%d" % (i, nm, cities[city[i]], score[i]))
    return b"\n".join(rows) + b"\n"


def write_csv_file(path: str, target_mb: int, seed: int = 3001, quoted_rows_per_chunk: int = 0):
    """C3-shaped file `name,city,score`: rows until the name column holds about target_mb MB.
    Assembled with numpy (a Python row loop would cost minutes at 40 M rows).  With
    quoted_rows_per_chunk > 0 every 2 M-row chunk is followed by that many rows whose name cell is
    quoted with an embedded ',' and a doubled '""' (RFC 4180).  -> (rows, file bytes)."""
    import os
    suffix = np.frombuffer(b",riverton,0421", dtype=np.uint8)
    ns = suffix.size
    col_bytes = 0
    total_rows = 0
    chunk = 0
    with open(path, "wb") as f:
        f.write(b"name,city,score\n")
        while col_bytes < target_mb * 1_000_000:
            rows = int(min(2_000_000, max(1000, (target_mb * 1_000_000 - col_bytes) // 14)))
            text, _ = company_names_joined(rows, seed=seed + 17 * chunk, upper=False)
            text = np.concatenate([text, np.frombuffer(b"\n", np.uint8)])
            nl = text == 10
            k = np.cumsum(nl, dtype=np.int64)
            dst = np.arange(text.size, dtype=np.int64) + ns * k
            out = np.empty(text.size + ns * int(k[-1]), dtype=np.uint8)
            out[dst] = text
            at = dst[nl] - ns
            out[(at[:, None] + np.arange(ns)[None, :]).ravel()] = np.tile(suffix, at.size)
            out.tofile(f)
            col_bytes += text.size
            total_rows += rows
            if quoted_rows_per_chunk:
                names = company_names(quoted_rows_per_chunk, seed=seed + 17 * chunk + 5, upper=False)
                blk = []
                for nm in names:
                    half = len(nm) // 2
                    blk.append(b'"' + nm[:half] + b', ""' + nm[half:] + b'"" co",lakeside,0007')
                    col_bytes += len(nm) + 11
                f.write(b"\n".join(blk) + b"\n")
                total_rows += quoted_rows_per_chunk
            chunk += 1
    return total_rows, os.path.getsize(path)
