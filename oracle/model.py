"""NumPy model of the DATA-PARALLEL decomposition of ORDER-E that the CUDA path implements.

TEST INFRASTRUCTURE ONLY (imported by tests/ only).  ``sa_oracle.c`` restates the reference's
recursive algorithm (engine.c:696-812); this file restates the *same order* as the sequence of
flat array passes the GPU runs, so that each CUDA stage has a CPU twin to be compared with:

  round 0     masked big-endian 8-byte key of every suffix, stable sort             (pack + onesweep)
  round j>0   groups still tied after D bytes, larger than 32, not ended by a terminator
              ('\\n' or 0xFF, engine.c:771-773,795-796) are re-sorted by the next `cb` bytes with a
              composite (dense group id, next bytes) key                             (refine rounds)
  finish      adjacent LCPs -> d*(r) = 1 + max over the 33-rank windows containing r of the min
              LCP inside the window = the depth at which r's bucket first has <= 32 members
              (engine.c:710); each such bucket is re-sorted by (W17(p + d*), p) where W17 is the
              17-byte window key of strncmp_128 (engine.c:664-677), unless d* >= L' (engine.c:706)
              or the shared d*-byte prefix holds a terminator                        (finish kernel)

``model_build`` must equal ``sao_build`` / the reference bit for bit (tests/test_model.py).
"""
from __future__ import annotations

import numpy as np

SMALL = 32
PAD = 512


def padded(text: np.ndarray) -> np.ndarray:
    t = np.zeros(text.size + PAD, dtype=np.uint8)
    t[: text.size] = text
    return t


def masked_words(T: np.ndarray, pos: np.ndarray, depth: int, cb: int) -> np.ndarray:
    """Big-endian u64 holding bytes [depth, depth+cb) of each suffix in its TOP cb bytes, bytes
    after the first terminator zeroed.  The caller guarantees no terminator before `depth`."""
    key = np.zeros(pos.size, dtype=np.uint64)
    alive = np.ones(pos.size, dtype=bool)
    for j in range(cb):
        b = T[pos.astype(np.int64) + depth + j].astype(np.uint64)
        b = np.where(alive, b, 0).astype(np.uint64)
        key |= b << np.uint64(8 * (7 - j))
        alive &= ~((b == 10) | (b == 255))
    return key


def has_terminator(key: np.ndarray, cb: int) -> np.ndarray:
    out = np.zeros(key.size, dtype=bool)
    for j in range(cb):
        b = (key >> np.uint64(8 * (7 - j))) & np.uint64(0xFF)
        out |= (b == 10) | (b == 255)
    return out


def common_bytes(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    """Number of equal leading bytes of two big-endian u64 arrays (8 when equal)."""
    x = a ^ b
    out = np.full(a.size, 8, dtype=np.int64)
    for j in range(7, -1, -1):
        byte = (x >> np.uint64(8 * (7 - j))) & np.uint64(0xFF)
        out = np.where(byte != 0, j, out)
    return out


def update_lcp(lcp, ranks, keys, same_group_as_prev, depth, cb, lprime):
    """lcp[ranks[i]] for pairs (i-1, i) inside one group, from the freshly sorted keys."""
    if ranks.size < 2:
        return
    cm = common_bytes(keys[1:], keys[:-1])
    eq = cm == 8
    # compare only the cb bytes that are part of the key (low bytes are zero in both)
    cm = np.minimum(cm, cb)
    val = depth + cm
    term = has_terminator(keys[1:], cb)
    val = np.where(eq & term, lprime, val)           # tied through a terminator: never refined
    val = np.minimum(val, lprime)
    sel = same_group_as_prev[1:]
    lcp[ranks[1:][sel]] = val[sel]


def refine_state(T, n, lprime, max_rounds=1000, trace=None):
    """Rounds 0..: returns (SA, lcp) with lcp[r] = LCP(rank r-1, rank r) as far as it was resolved."""
    pos = np.arange(n, dtype=np.uint32)
    cb = min(8, lprime)
    key = masked_words(T, pos, 0, cb)
    order = np.argsort(key, kind="stable")
    SA = pos[order]
    k = key[order]
    lcp = np.zeros(n, dtype=np.int64)
    update_lcp(lcp, np.arange(n), k, np.ones(n, dtype=bool), 0, cb, lprime)
    depth = cb
    if trace is not None:
        trace.append(dict(round=0, depth=depth, SA=SA.copy(), lcp=lcp.copy(), active=n))
    rnd = 0
    while depth < lprime and rnd < max_rounds:
        rnd += 1
        open_tie = lcp == depth                      # pair (r-1, r) tied on `depth` bytes, no terminator
        open_tie[0] = False
        # group = a rank that is not an open tie followed by >= 1 open ties
        nxt = np.zeros(n, dtype=bool)
        nxt[:-1] = open_tie[1:]
        head = ~open_tie & nxt
        member = open_tie | head
        # group sizes
        head_idx = np.flatnonzero(head)
        if head_idx.size == 0:
            break
        # end of each group: last consecutive open tie
        not_member_after = np.flatnonzero(~nxt)      # ranks whose successor is not an open tie
        ends = not_member_after[np.searchsorted(not_member_after, head_idx)]
        sizes = ends - head_idx + 1
        big = sizes > SMALL
        if not big.any():
            break
        gh = head_idx[big]
        ge = ends[big]
        ngroups = gh.size
        gbits = max(1, int(ngroups - 1).bit_length())
        cb = min(8, (64 - gbits) // 8, lprime - depth)
        assert cb >= 1
        slots = np.concatenate([np.arange(h, e + 1) for h, e in zip(gh, ge)])
        gid = np.repeat(np.arange(ngroups), ge - gh + 1)
        p = SA[slots]
        key = masked_words(T, p, depth, cb)
        order = np.lexsort((key, gid))               # stable: ties keep ascending position
        SA[slots] = p[order]
        k = key[order]
        g2 = gid[order]
        same = np.ones(slots.size, dtype=bool)
        same[0] = False
        same[1:] = g2[1:] == g2[:-1]
        update_lcp(lcp, slots, k, same, depth, cb, lprime)
        depth += cb
        if trace is not None:
            trace.append(dict(round=rnd, depth=depth, SA=SA.copy(), lcp=lcp.copy(), active=slots.size,
                              ngroups=ngroups, gbits=gbits, cb=cb))
    return SA, lcp


def dstar_from_lcp(lcp: np.ndarray, n: int) -> np.ndarray:
    """d*(r) = 1 + max over windows [s, s+32] containing r of min(lcp[s+1 .. s+32]); 0 if n <= 32."""
    if n <= SMALL:
        return np.zeros(n, dtype=np.int64)
    nwin = n - SMALL                                 # window starts s = 0 .. n-33
    m = np.full(nwin, np.iinfo(np.int64).max)
    for j in range(1, SMALL + 1):
        m = np.minimum(m, lcp[j: j + nwin])
    M = np.full(n, -1, dtype=np.int64)
    for j in range(0, SMALL + 1):                    # r = s + j
        seg = M[j: j + nwin]
        np.maximum(seg, m, out=seg)
    return M + 1


def w17_key(T: np.ndarray, p: int, lprime: int) -> bytes:
    """17-byte sort key equivalent to strncmp_128's compare (engine.c:664-677): the 16-byte window
    with everything after its first NUL zeroed, then the raw 17th byte; only the first
    min(L', 17) bytes take part (the idx < max_depth gate)."""
    w = bytearray(T[p: p + 17].tobytes())
    for i in range(16):
        if w[i] == 0:
            for j in range(i, 16):
                w[j] = 0
            break
    return bytes(w[: min(lprime, 17)])


def finish(T, SA, lcp, n, lprime, trace=None):
    dstar = dstar_from_lcp(lcp, n)
    out = SA.copy()
    r = 0
    runs = []
    while r < n:
        d = int(dstar[r])
        e = r
        while e + 1 < n and lcp[e + 1] >= d:
            e += 1
        size = e - r + 1
        assert size <= SMALL or d >= lprime, (r, e, d)
        if d < lprime and size >= 2:
            p0 = int(SA[r])
            pref = T[p0: p0 + d]
            if not ((pref == 10) | (pref == 255)).any():
                members = [int(x) for x in SA[r: e + 1]]
                members.sort(key=lambda p: (w17_key(T, p + d, lprime), p))
                out[r: e + 1] = members
                runs.append((r, e, d))
        r = e + 1
    if trace is not None:
        trace.append(dict(stage="finish", dstar=dstar, runs=runs))
    return out


def model_build(text, L: int, trace=None) -> np.ndarray:
    text = np.ascontiguousarray(text, dtype=np.uint8)
    n = text.size
    if n == 0:
        return np.zeros(0, dtype=np.uint32)
    lprime = min(L, n)
    T = padded(text)
    if lprime == 0:
        return np.arange(n, dtype=np.uint32)
    SA, lcp = refine_state(T, n, lprime, trace=trace)
    return finish(T, SA, lcp, n, lprime, trace=trace).astype(np.uint32)


# ------------------------------------------------------------------------------ CSV column scan (csv_scan.cu)
CSV_IRREGULAR_NUL = 1
CSV_IRREGULAR_QUOTE = 2


def csv_scan_model(data: bytes, wanted: int, num_columns: int):
    """NumPy twin of the device column scanner (csrc/csv_scan.cu: quote parity -> unquoted delimiter
    counts -> emit): the reference's byte state machine (engine.c:522-605) restated as prefix sums,
    with the same irregularity flags.  -> (flags, text u8, map u32 with the quoted flag in bit 31).
    For flags == 0 the streams must equal sao_csv_scan's (tests/test_model.py)."""
    f = np.frombuffer(data, np.uint8)
    size = f.size
    flags = CSV_IRREGULAR_NUL if (f == 0).any() else 0
    q = f == 34
    inq_before = ((np.cumsum(q) - q) & 1).astype(bool)                 # in quotes in front of byte p
    draw = (f == 44) | (f == 10)
    d = draw & ~inq_before                                             # delimiters outside quotes
    col = (np.cumsum(d) - d) % num_columns                             # a delimiter belongs to the cell it ends
    prev = np.concatenate([[44], f[:-1]]) if size else f               # the file start behaves like a cell start
    nxt = np.concatenate([f[1:], [0]]) if size else f
    qprev, qnext = prev == 34, nxt == 34
    opening, closing = q & ~inq_before, q & inq_before
    if (opening & ~((prev == 44) | (prev == 10) | qprev)).any():
        flags |= CSV_IRREGULAR_QUOTE
    dnext = (nxt == 44) | (nxt == 10)
    if size:
        dnext[-1] = True                                               # the byte behind the last one counts as a delimiter
    if (closing & ~(dnext | qnext)).any():
        flags |= CSV_IRREGULAR_QUOTE
    if int(q.sum()) & 1:
        flags |= CSV_IRREGULAR_QUOTE                                   # unterminated quote
    escaped = closing & qnext                                          # first quote of a "" pair inside quotes
    emit = (col == wanted) & (~q | escaped)
    p = np.flatnonzero(emit)
    c = f[p].copy()
    up = (c >= 65) & (c <= 90) & ~d[p]
    c[up] += 32
    c[d[p]] = 10                                                       # the cell's delimiter -> row sentinel
    qf = np.where(d[p], qprev[p], inq_before[p]).astype(np.uint32)
    mp = p.astype(np.uint32) | (qf << np.uint32(31))
    if size:                                                           # end of the file inside the wanted cell
        last = int(f[-1])
        last_is_delim = last in (44, 10) and not inq_before[-1]
        if not last_is_delim and int(d.sum()) % num_columns == wanted:
            c = np.concatenate([c, [10]]).astype(np.uint8)
            mp = np.concatenate([mp, [np.uint32(size) | (np.uint32(1 << 31) if last == 34 else np.uint32(0))]]).astype(np.uint32)
    return flags, c.astype(np.uint8), mp.astype(np.uint32)


# ------------------------------------------------------------------------------ finish, per-rank rules (finish.cu)
def finish_flat(T, SA, lcp, n, lprime, lo=0, hi=None):
    """finish() the way finish_kernel decides it -- per rank, with no walk from rank 0: rank j heads a
    bucket iff lcp[j] < d*(j); a bucket is re-sorted iff d* < L', it has two members or more and its
    shared prefix holds no terminator.  Only ranks in [lo, hi) are placed (the rest of `out` keeps the
    input order): with SA / lcp = [left halo | own shard | right halo] this is the sharded finish
    (sa_shard.cu), which must reproduce the slice of the global result when the halos are >= 96 ranks
    (or reach the ends of the array)."""
    hi = n if hi is None else hi
    dstar = dstar_from_lcp(lcp, n)
    out = SA.copy()
    head = np.ones(n, dtype=bool)
    head[1:] = lcp[1:n] < dstar[1:]
    starts = np.flatnonzero(head)
    ends = np.concatenate([starts[1:], [n]])
    for r, e1 in zip(starts.tolist(), ends.tolist()):
        if e1 <= lo or r >= hi:
            continue
        d, size = int(dstar[r]), e1 - r
        if d >= lprime or size < 2:
            continue
        assert size <= SMALL, (r, e1, d)
        p0 = int(SA[r])
        pref = T[p0: p0 + d]
        if ((pref == 10) | (pref == 255)).any():
            continue
        members = sorted((int(x) for x in SA[r: e1]), key=lambda p: (w17_key(T, p + d, lprime), p))
        for x, p in zip(range(r, e1), members):
            if lo <= x < hi:
                out[x] = p
    return out
