"""One suffix array over several GPUs: host-side driver of the sample-sort build.

Every rank holds the whole text (replicated) and one ``sa_index``.  The phases and the
collective between each pair of them:

  pack      keys + positions of this rank's slice of text positions, 65536-bin key histogram
            -- all-reduce(SUM) of the histogram (NCCL) --
  split     splitters from the global histogram (equal keys never straddle ranks), stable
            split of the local records by destination rank
            -- all-to-all of the (key, position) records (NCCL over NVLink) --
  sort      LSD radix sort + tie refinement of the received records (sa_shard_sort)
            -- all-gather of each rank's first/last 128 (position, lcp) records (tiny) --
  finish    <= 32 bucket finish with a 128-record halo from each neighbour (sa_shard_finish)
            -- all-gather of the finished slices --
  adopt     every rank installs the complete SA and answers its share of a query batch

``build_distributed`` runs one rank of this under ``torch.distributed`` (NCCL on GPUs);
``build_emulated`` runs G ranks in lock step inside ONE process on one GPU with the collectives
done by plain tensor copies -- it exists so the single-GPU test tier can check the sharded
pipeline bit for bit against the oracle.  The reference has no multi-device path at all; its
partitioning into independent indexes is suffix_array.pyx:148-180.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib as S

EDGE = 128           # halo records fetched from each side (>= 96 needed by the finish kernel)
NBINS = 65536


# ----------------------------------------------------------------------------- host logic (NumPy)
def slice_bounds(n: int, world: int) -> list[int]:
    """Text-position slices per rank: equal parts, starts rounded down to a multiple of 16."""
    b = [(g * n // world) & ~15 for g in range(world)]
    b[0] = 0
    return b + [n]


def choose_bin2rank(hist: np.ndarray, world: int) -> np.ndarray:
    """Destination rank of every top-16-bit key bin: bins are assigned in order so that rank g
    gets the bins whose first record falls into the g-th equal share of the records."""
    hist = np.asarray(hist, dtype=np.int64)
    total = int(hist.sum())
    before = np.cumsum(hist) - hist
    if total == 0:
        return np.zeros(NBINS, dtype=np.uint8)
    r = (before * world) // total
    return np.minimum(r, world - 1).astype(np.uint8)


def _lower(b: np.ndarray) -> np.ndarray:
    b = b.copy()
    up = (b >= 65) & (b <= 90)
    b[up] += 32
    return b


def boundary_lcp(text: np.ndarray, p: int, q: int, lprime: int, lowercase: bool) -> int:
    """LCP (bytes, capped at L') of the masked prefixes of suffixes p and q, as the device
    computes it for records adjacent in prefix order: equal through a terminator means tied for
    good (L')."""
    n = text.size
    a = np.zeros(lprime, dtype=np.uint8)
    b = np.zeros(lprime, dtype=np.uint8)
    ta = text[p:min(n, p + lprime)]
    tb = text[q:min(n, q + lprime)]
    a[:ta.size] = ta
    b[:tb.size] = tb
    if lowercase:
        a, b = _lower(a), _lower(b)
    for i in range(lprime):
        if a[i] != b[i]:
            return i
        if a[i] == 10 or a[i] == 255:
            return lprime
    return lprime


def assemble_halos(edges: list[dict], rank: int, text: np.ndarray, lprime: int, lowercase: bool):
    """edges[g] = {"m", "head_pos", "head_lcp", "tail_pos", "tail_lcp"} of every rank.
    -> (left_pos, left_lcp, lcp_first, right_pos, right_lcp) for `rank` (global order)."""
    me = edges[rank]
    left_pos: list[np.ndarray] = []
    left_lcp: list[np.ndarray] = []
    need = EDGE
    nxt_first = int(me["head_pos"][0])          # first record of the chunk to the right of the gap
    lcp_first = 0
    have_left = False
    g = rank - 1
    while g >= 0 and need > 0:
        e = edges[g]
        k = min(need, int(e["tail_pos"].size))
        if k > 0:
            pos = e["tail_pos"][-k:].astype(np.uint32)
            lcp = e["tail_lcp"][-k:].astype(np.uint16).copy()
            junction = boundary_lcp(text, int(pos[-1]), nxt_first, lprime, lowercase)
            if not have_left:
                lcp_first = junction
                have_left = True
            else:
                left_lcp[0][0] = junction       # first record of the chunk collected before
            left_pos.insert(0, pos)
            left_lcp.insert(0, lcp)
            nxt_first = int(pos[0])
            need -= k
            if k < int(e["m"]):                 # this rank still has records in front: halo is full
                break
        g -= 1
    right_pos: list[np.ndarray] = []
    right_lcp: list[np.ndarray] = []
    need = EDGE
    prev_last = int(me["tail_pos"][-1])
    g = rank + 1
    while g < len(edges) and need > 0:
        e = edges[g]
        k = min(need, int(e["head_pos"].size))
        if k > 0:
            pos = e["head_pos"][:k].astype(np.uint32)
            lcp = e["head_lcp"][:k].astype(np.uint16).copy()
            lcp[0] = boundary_lcp(text, prev_last, int(pos[0]), lprime, lowercase)
            right_pos.append(pos)
            right_lcp.append(lcp)
            prev_last = int(pos[-1])
            need -= k
            if k < int(e["m"]):
                break
        g += 1
    cat = lambda xs, dt: np.concatenate(xs).astype(dt) if xs else np.zeros(0, dt)   # noqa: E731
    return (cat(left_pos, np.uint32), cat(left_lcp, np.uint16), int(lcp_first),
            cat(right_pos, np.uint32), cat(right_lcp, np.uint16))


class DeviceText:
    """The text of a shard build that only exists on the device (a scanned CSV column): `size` and
    slicing, which is all boundary_lcp needs -- a window is fetched when it is asked for."""

    def __init__(self, idx, n: int):
        self._idx, self.size = idx, int(n)

    def __getitem__(self, sl):
        a, b, step = sl.indices(self.size)
        assert step == 1
        out = np.empty(max(b - a, 0), dtype=np.uint8)
        if out.size:
            S.check(S.lib().sa_shard_copy_text(self._idx._h, a, out.size, out.ctypes.data_as(S._vp)))
        return out

    def whole(self) -> np.ndarray:
        return self[0:self.size]


# ----------------------------------------------------------------------------- one rank
class ShardBuilder:
    """The per-rank phases.  Buffers that cross a collective are torch CUDA tensors."""

    def __init__(self, text, max_suffix_length: int, rank: int, world: int, device: int,
                 lowercase: bool = False, timing: bool = False, device_text=None, csv=None):
        """text: the whole text on the host (also used for the few boundary LCPs); device_text: an
        optional torch CUDA uint8 tensor holding the same bytes, to skip the host-to-device copy.
        csv = (path, column): the text is that column of a CSV file instead (every rank scans the
        file itself; the finished array holds file offsets, as after sa_build_csv)."""
        import torch
        self.torch = torch
        self.rank, self.world, self.device = rank, world, device
        self.L = int(max_suffix_length)
        self.dev = torch.device("cuda", device)
        self.idx = S.Index(self.L, device=device)
        if csv is not None:
            self.begin_csv(csv[0], csv[1], timing)
        else:
            self.begin(text, lowercase, timing, device_text)

    def begin(self, text: np.ndarray, lowercase: bool = False, timing: bool = False, device_text=None):
        """(Re)start a build on this rank's index; scratch from an earlier build is reused."""
        self.text = np.ascontiguousarray(text, dtype=np.uint8)
        self.n = int(self.text.size)
        self.lowercase = lowercase
        self.lprime = min(self.L, self.n)
        flags = (S.SA_FLAG_LOWERCASE if lowercase else 0) | (S.SA_FLAG_TIMING if timing else 0)
        if device_text is not None:
            assert device_text.numel() == self.n and device_text.is_cuda
            S.check(S.lib().sa_shard_begin(self.idx._h, S._vp(device_text.data_ptr()), self.n,
                                           flags | S.SA_FLAG_DEVICE_TEXT))
        else:
            S.check(S.lib().sa_shard_begin(self.idx._h, self.text.ctypes.data_as(S._vp), self.n, flags))
        b = slice_bounds(self.n, self.world)
        self.pos_begin, self.pos_end = b[self.rank], b[self.rank + 1]
        self.m_local = self.pos_end - self.pos_begin

    def begin_csv(self, path: str, column: str, timing: bool = False):
        """(Re)start a build whose text is a CSV column (sa_shard_begin_csv)."""
        S.check(S.lib().sa_shard_begin_csv(self.idx._h, path.encode(), column.encode(), S.SA_FLAG_TIMING if timing else 0))
        self.n = int(S.lib().sa_shard_text_size(self.idx._h))
        self.text = DeviceText(self.idx, self.n)               # windows are fetched for the boundary LCPs only
        self.lowercase = False                                 # the column scan lower-cases (engine.c:559,581,593)
        self.lprime = min(self.L, self.n)
        b = slice_bounds(self.n, self.world)
        self.pos_begin, self.pos_end = b[self.rank], b[self.rank + 1]
        self.m_local = self.pos_end - self.pos_begin

    def _p(self, t):
        return S._vp(t.data_ptr()) if t.numel() else None

    def pack(self):
        torch = self.torch
        self.keys = torch.empty(max(self.m_local, 1), dtype=torch.int64, device=self.dev)
        self.pos = torch.empty(max(self.m_local, 1), dtype=torch.int32, device=self.dev)
        self.hist = torch.zeros(NBINS, dtype=torch.int32, device=self.dev)
        torch.cuda.synchronize(self.dev)
        S.check(S.lib().sa_shard_pack(self.idx._h, self.pos_begin, self.pos_end, self._p(self.keys), self._p(self.pos),
                                      self._p(self.hist)))
        return self.hist

    def split(self, global_hist_np: np.ndarray):
        torch = self.torch
        self.bin2rank = choose_bin2rank(global_hist_np, self.world)
        self.send_keys = torch.empty_like(self.keys)
        self.send_pos = torch.empty_like(self.pos)
        counts = np.zeros(self.world, dtype=np.uint64)
        torch.cuda.synchronize(self.dev)
        S.check(S.lib().sa_shard_split(self.idx._h, self._p(self.keys), self._p(self.pos), self.m_local,
                                       self.bin2rank.ctypes.data_as(S._vp), self.world, self._p(self.send_keys),
                                       self._p(self.send_pos), counts.ctypes.data_as(S._vp)))
        self.keys = self.pos = None
        self.send_counts = counts.astype(np.int64)
        return self.send_counts

    # ---- fused split + exchange (peer stores through CUDA IPC)
    def count(self, global_hist_np: np.ndarray):
        self.bin2rank = choose_bin2rank(global_hist_np, self.world)
        counts = np.zeros(self.world, dtype=np.uint64)
        self.torch.cuda.synchronize(self.dev)
        S.check(S.lib().sa_shard_count(self.idx._h, self._p(self.keys), self.m_local,
                                       self.bin2rank.ctypes.data_as(S._vp), self.world, counts.ctypes.data_as(S._vp)))
        self.send_counts = counts.astype(np.int64)
        return self.send_counts

    def recv_prepare(self, m_recv: int) -> np.ndarray:
        """-> 128 bytes: the IPC handles of this rank's receive (= sort) buffers."""
        self.m = int(m_recv)
        h = np.zeros(128, dtype=np.uint8)
        S.check(S.lib().sa_shard_recv_prepare(self.idx._h, self.m, h[:64].ctypes.data_as(S._vp),
                                              h[64:].ctypes.data_as(S._vp)))
        return h

    def scatter_p2p(self, all_handles: np.ndarray, counts_matrix: np.ndarray):
        """counts_matrix[src, dst]; this rank's run starts in rank d's buffers after the runs of lower ranks."""
        dst_off = counts_matrix[: self.rank, :].sum(axis=0).astype(np.uint64)
        handles = np.ascontiguousarray(all_handles, dtype=np.uint8).reshape(-1)
        S.check(S.lib().sa_shard_scatter_p2p(self.idx._h, self._p(self.keys), self._p(self.pos), self.m_local,
                                             self.world, self.rank, handles.ctypes.data_as(S._vp),
                                             dst_off.ctypes.data_as(S._vp)))
        self.keys = self.pos = None

    def sort_recv(self):
        S.check(S.lib().sa_shard_sort_recv(self.idx._h, self.m))

    def sort(self, recv_keys, recv_pos, m: int):
        self.m = int(m)
        self.torch.cuda.synchronize(self.dev)
        S.check(S.lib().sa_shard_sort(self.idx._h, self._p(recv_keys), self._p(recv_pos), self.m))
        self.send_keys = self.send_pos = None

    def edges(self) -> dict:
        k = min(EDGE, self.m)
        hp, tp = np.zeros(k, np.uint32), np.zeros(k, np.uint32)
        hl, tl = np.zeros(k, np.uint16), np.zeros(k, np.uint16)
        if k:
            S.check(S.lib().sa_shard_edges(self.idx._h, EDGE, hp.ctypes.data_as(S._vp), hl.ctypes.data_as(S._vp),
                                           tp.ctypes.data_as(S._vp), tl.ctypes.data_as(S._vp)))
        return {"m": self.m, "head_pos": hp, "head_lcp": hl, "tail_pos": tp, "tail_lcp": tl}

    def finish(self, all_edges: list[dict]):
        torch = self.torch
        out = torch.empty(max(self.m, 1), dtype=torch.int32, device=self.dev)
        if self.m:
            lp, ll, lf, rp, rl = assemble_halos(all_edges, self.rank, self.text, self.lprime, self.lowercase)
            vp = lambda a: a.ctypes.data_as(S._vp) if a.size else None   # noqa: E731
            torch.cuda.synchronize(self.dev)
            S.check(S.lib().sa_shard_finish(self.idx._h, vp(lp), vp(ll), lp.size, lf, vp(rp), vp(rl), rp.size,
                                            self._p(out)))
        self.slice = out[: self.m]
        return self.slice

    # ---- fused finish + replicate (peer stores through CUDA IPC)
    def final_prepare(self) -> np.ndarray:
        """-> 64 bytes: the IPC handle of this rank's final array."""
        h = np.zeros(64, dtype=np.uint8)
        S.check(S.lib().sa_shard_final_prepare(self.idx._h, h.ctypes.data_as(S._vp)))
        return h

    def finish_bcast(self, all_edges: list[dict], all_handles: np.ndarray):
        """finish this shard and store its slice into every rank's final array."""
        goff = int(sum(int(e["m"]) for e in all_edges[: self.rank]))
        handles = np.ascontiguousarray(all_handles, dtype=np.uint8).reshape(-1)
        z32, z16 = np.zeros(0, np.uint32), np.zeros(0, np.uint16)
        lp, ll, lf, rp, rl = (z32, z16, 0, z32, z16)
        if self.m:
            lp, ll, lf, rp, rl = assemble_halos(all_edges, self.rank, self.text, self.lprime, self.lowercase)
        vp = lambda a: a.ctypes.data_as(S._vp) if a.size else None   # noqa: E731
        self.torch.cuda.synchronize(self.dev)
        S.check(S.lib().sa_shard_finish_bcast(self.idx._h, vp(lp), vp(ll), lp.size, lf, vp(rp), vp(rl), rp.size,
                                              self.world, self.rank, handles.ctypes.data_as(S._vp), goff))

    def mark_built(self):
        S.check(S.lib().sa_shard_mark_built(self.idx._h))

    def adopt(self, full_sa):
        self.torch.cuda.synchronize(self.dev)
        S.check(S.lib().sa_adopt_sa(self.idx._h, self._p(full_sa), self.n))

    def adopt_parts(self, gathered, stride: int, counts):
        c = np.asarray(counts, dtype=np.uint64)
        self.torch.cuda.synchronize(self.dev)
        S.check(S.lib().sa_adopt_sa_parts(self.idx._h, self._p(gathered), int(stride), c.ctypes.data_as(S._vp), int(c.size)))

    def close(self):
        self.idx.close()


def _edges_to_array(e: dict) -> np.ndarray:
    """Fixed-size int64 row for the all-gather: [m, k, head_pos.., head_lcp.., tail_pos.., tail_lcp..]."""
    row = np.zeros(2 + 4 * EDGE, dtype=np.int64)
    k = int(e["head_pos"].size)
    row[0], row[1] = e["m"], k
    row[2:2 + k] = e["head_pos"]
    row[2 + EDGE:2 + EDGE + k] = e["head_lcp"]
    row[2 + 2 * EDGE:2 + 2 * EDGE + k] = e["tail_pos"]
    row[2 + 3 * EDGE:2 + 3 * EDGE + k] = e["tail_lcp"]
    return row


def _array_to_edges(row: np.ndarray) -> dict:
    m, k = int(row[0]), int(row[1])
    return {"m": m, "head_pos": row[2:2 + k].astype(np.uint32), "head_lcp": row[2 + EDGE:2 + EDGE + k].astype(np.uint16),
            "tail_pos": row[2 + 2 * EDGE:2 + 2 * EDGE + k].astype(np.uint32),
            "tail_lcp": row[2 + 3 * EDGE:2 + 3 * EDGE + k].astype(np.uint16)}


# ----------------------------------------------------------------------------- drivers
def build_emulated(text, max_suffix_length: int, world: int, device: int = 0, lowercase: bool = False,
                   adopt: bool = True, exchange: str = "p2p", replicate: str = "allgather", csv=None):
    """G ranks in lock step inside one process on ONE GPU (collectives = tensor copies).
    -> (full array of text positions as a torch int32 CUDA tensor, list of ShardBuilder).  Test tier only."""
    import torch
    if csv is None:
        text = np.ascontiguousarray(np.frombuffer(text, np.uint8) if isinstance(text, (bytes, bytearray)) else text,
                                    dtype=np.uint8)
    ranks = [ShardBuilder(text, max_suffix_length, g, world, device, lowercase, csv=csv) for g in range(world)]
    hist = sum(r.pack().to(torch.int64) for r in ranks)                       # all-reduce
    hist_np = hist.cpu().numpy()
    if exchange == "p2p":
        for r in ranks:
            r.idx.debug_set("same_process", 1)
        counts = np.stack([r.count(hist_np) for r in ranks])                   # all-gather of the counts
        handles = np.stack([r.recv_prepare(int(counts[:, d].sum())) for d, r in enumerate(ranks)])   # all-gather
        for r in ranks:
            r.scatter_p2p(handles, counts)
        for r in ranks:                                                        # barrier
            r.sort_recv()
        edges = [_array_to_edges(_edges_to_array(r.edges())) for r in ranks]
        if replicate == "fused":
            fh = np.stack([r.final_prepare() for r in ranks])                   # all-gather of the handles
            for r in ranks:
                r.finish_bcast(edges, fh)
            for r in ranks:                                                    # barrier
                r.mark_built()
            full = torch.from_numpy(ranks[0].idx.text_positions().astype(np.int32)).to(ranks[0].dev)
            return full, ranks
        slices = [r.finish(edges) for r in ranks]
        full = torch.cat(slices)
        if adopt:
            for r in ranks:
                r.adopt(full)
        return full, ranks
    counts = np.stack([r.split(hist_np) for r in ranks])                       # [src, dst]
    for d, r in enumerate(ranks):                                              # all-to-all
        parts_k, parts_p = [], []
        for s_, src in enumerate(ranks):
            off = int(counts[s_, :d].sum())
            c = int(counts[s_, d])
            parts_k.append(src.send_keys[off:off + c])
            parts_p.append(src.send_pos[off:off + c])
        rk = torch.cat(parts_k) if parts_k else torch.empty(0, dtype=torch.int64, device=r.dev)
        rp = torch.cat(parts_p) if parts_p else torch.empty(0, dtype=torch.int32, device=r.dev)
        r._recv = (rk, rp, int(counts[:, d].sum()))
    for r in ranks:
        r.sort(*r._recv)
        r._recv = None
    edges = [_array_to_edges(_edges_to_array(r.edges())) for r in ranks]       # all-gather
    slices = [r.finish(edges) for r in ranks]
    full = torch.cat(slices)                                                   # all-gather
    if adopt:
        for r in ranks:
            r.adopt(full)
    return full, ranks


def build_distributed(text, max_suffix_length: int, device: int, lowercase: bool = False, timing: bool = False,
                      adopt: bool = True, group=None, device_text=None, builder=None, exchange: str = "p2p",
                      replicate: str = "fused", csv=None):
    """This process's rank of the sample-sort build under torch.distributed (NCCL).
    csv = (path, column): index that column of a CSV file instead of `text`.
    -> (ShardBuilder with the complete SA adopted, dict of per-phase wall times in ms)."""
    import time
    import torch
    import torch.distributed as dist
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    if csv is None:
        text = np.ascontiguousarray(np.frombuffer(text, np.uint8) if isinstance(text, (bytes, bytearray)) else text,
                                    dtype=np.uint8)
    t = {}
    t0 = time.perf_counter()

    def lap(name):
        nonlocal t0
        if torch.cuda.is_available():          # (the host-logic tests drive this function with a CPU stand-in over gloo)
            torch.cuda.synchronize(device)
        now = time.perf_counter()
        t[name] = (now - t0) * 1e3
        t0 = now

    if builder is None:
        r = ShardBuilder(text, max_suffix_length, rank, world, device, lowercase, timing, device_text, csv=csv)
    elif csv is not None:
        r = builder
        r.begin_csv(csv[0], csv[1], timing)
    else:
        r = builder
        r.begin(text, lowercase, timing, device_text)
    lap("begin")
    hist = r.pack().to(torch.int64)
    dist.all_reduce(hist, op=dist.ReduceOp.SUM, group=group)
    hist_np = hist.cpu().numpy()
    lap("pack+hist")
    if exchange == "p2p":
        send_counts = r.count(hist_np)
        sc = torch.from_numpy(send_counts.copy()).to(r.dev)
        rows_c = [torch.empty_like(sc) for _ in range(world)]
        dist.all_gather(rows_c, sc, group=group)
        cm = np.stack([x.cpu().numpy() for x in rows_c])                       # [src, dst]
        m = int(cm[:, rank].sum())
        hnd = torch.from_numpy(r.recv_prepare(m)).to(r.dev)
        rows_h = [torch.empty_like(hnd) for _ in range(world)]
        dist.all_gather(rows_h, hnd, group=group)                             # also: every rank's buffers are idle now
        handles = np.stack([x.cpu().numpy() for x in rows_h])
        lap("split")
        r.scatter_p2p(handles, cm)
        dist.barrier(group=group)                                              # all peer stores have landed
        lap("exchange")
        r.sort_recv()
        lap("sort+refine")
        r.exchange_bytes = int((send_counts.sum() - send_counts[rank]) * 12)
        return _finish_distributed(r, t, lap, adopt, group, replicate)
    send_counts = r.split(hist_np)
    sc = torch.from_numpy(send_counts.copy()).to(r.dev)
    rc = torch.empty_like(sc)
    dist.all_to_all_single(rc, sc, group=group)
    recv_counts = rc.cpu().numpy()
    m = int(recv_counts.sum())
    lap("split")
    recv_keys = torch.empty(max(m, 1), dtype=torch.int64, device=r.dev)
    recv_pos = torch.empty(max(m, 1), dtype=torch.int32, device=r.dev)
    ins, outs = send_counts.tolist(), recv_counts.tolist()
    dist.all_to_all_single(recv_keys[:m], r.send_keys[:r.m_local], outs, ins, group=group)
    dist.all_to_all_single(recv_pos[:m], r.send_pos[:r.m_local], outs, ins, group=group)
    lap("exchange")
    r.sort(recv_keys, recv_pos, m)
    del recv_keys, recv_pos
    lap("sort+refine")
    r.exchange_bytes = int((send_counts.sum() - send_counts[rank]) * 12)
    return _finish_distributed(r, t, lap, adopt, group, replicate)


def _finish_distributed(r, t, lap, adopt, group, replicate="fused"):
    import torch
    import torch.distributed as dist
    world = r.world
    fused = adopt and replicate == "fused"
    row = _edges_to_array(r.edges())
    if fused:                                   # the final array's IPC handle rides along (8 int64 words)
        row = np.concatenate([row, r.final_prepare().view(np.int64)])
    mine = torch.from_numpy(row).to(r.dev)
    rows = torch.empty(world * mine.numel(), dtype=mine.dtype, device=r.dev)
    dist.all_gather_into_tensor(rows, mine, group=group)
    rows_np = rows.cpu().numpy().reshape(world, -1)
    edges = [_array_to_edges(rows_np[g]) for g in range(world)]
    lap("edges")
    if fused:
        handles = np.ascontiguousarray(rows_np[:, -8:]).view(np.uint8).reshape(world, 64)
        r.finish_bcast(edges, handles)
        dist.barrier(group=group)               # every rank's stores into every final array have completed
        r.mark_built()
        r.replicated_bytes = int(r.m) * 4 * (world - 1)
        lap("finish+replicate")
        return r, t
    sl = r.finish(edges)
    lap("finish")
    if adopt:
        ms = [int(e["m"]) for e in edges]
        mmax = max(max(ms), 1)
        if r.m == mmax:
            padded = sl
        else:
            padded = torch.zeros(mmax, dtype=torch.int32, device=r.dev)
            padded[: r.m] = sl
        gathered = torch.empty(world * mmax, dtype=torch.int32, device=r.dev)
        dist.all_gather_into_tensor(gathered, padded, group=group)
        lap("gather")
        r.adopt_parts(gathered, mmax, ms)
        del gathered, padded
        lap("adopt")
    return r, t
