"""CPU-only checks of the host side: the package imports without a GPU, fails loudly when used,
the bench's reference arm emits the contract's JSON line, and the multi-process (gloo,
world_size 2) aggregation the bench uses for N > 1 gives max-over-ranks times / summed units."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _cuda():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_cuda(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    from suffixarray_b200 import SuffixArray
    from suffixarray_b200._lib import SaCudaError
    with pytest.raises(SaCudaError):
        SuffixArray(documents=["a", "b"], max_suffix_length=8)


def test_reference_arm_json_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "small",
                        "--steps", "1", "--warmup", "0"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["value"] > 0 and line["queries_per_s"] > 0
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["cpu_baseline"]["kind"] in ("reference", "port")


WORKER = r"""
import os, sys, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
import bench
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
# each rank owns one independent partition (suffix_array.pyx:148-180): seeds differ, sizes match
t, qb, qo, low = bench.make_workload("small", rank)
sums = torch.tensor([10.0 + rank, 3.0 - rank], dtype=torch.float64)
counts = torch.tensor([float(t.size), float(qo.size - 1)], dtype=torch.float64)
dist.all_reduce(sums, op=dist.ReduceOp.MAX)
dist.all_reduce(counts, op=dist.ReduceOp.SUM)
digest = torch.tensor([float(int(t[:4096].astype('int64').sum()))], dtype=torch.float64)
both = [torch.zeros_like(digest) for _ in range(world)]
dist.all_gather(both, digest)
if rank == 0:
    import json
    print("RESULT" + json.dumps([sums.tolist(), counts.tolist(), [b.item() for b in both]]))
dist.destroy_process_group()
"""


def test_gloo_world2_aggregation(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29731", str(script), ROOT],
                       capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stderr[-3000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT")][0]
    sums, counts, digests = json.loads(line[len("RESULT"):])
    assert sums == [11.0, 3.0]                       # max over ranks
    assert counts == [8_000_000.0, 100_000.0]        # units of all ranks
    assert digests[0] != digests[1]                  # distinct partitions


def test_csv_row_table_definition(tmp_path):
    """The row table the device builds (csv_rows_kernel -> sa_copy_row_starts, records cut by
    sa_query_records) is checked on the GPU against oracle_csv_row_starts; here that definition --
    rows end at '\n' outside quotes, header = row 0 -- is checked against Python's csv module on a file
    with raw newlines inside quoted cells, ragged rows and CRLF."""
    import csv
    import io
    from oracle import pyoracle as O
    rows = [["id", "name", "note"], ["1", "plain", "x"], ["2", "two\nlines", 'say "hi"'], ["3", "a,b", "three\n\nlines"],
            ["4"], ["5", "crlf\r\ninside", "y", "extra"], ["6", "last", ""]]
    for term, trailing in (("\n", True), ("\n", False), ("\r\n", True)):
        buf = io.StringIO()
        csv.writer(buf, lineterminator=term).writerows(rows)
        data = (buf.getvalue() if trailing else buf.getvalue()[: -len(term)]).encode()
        starts = O.oracle_csv_row_starts(data)
        nrows = starts.size - 1 - (1 if trailing else 0)          # a trailing '\n' opens one empty last row
        assert nrows == len(rows)
        got = []
        for d in range(nrows):
            end = int(starts[d + 1]) - (1 if d + 1 < starts.size - 1 else 0)
            line = data[int(starts[d]):end].decode().rstrip("\r")
            got.append(next(csv.reader([line])))
        assert got == rows


SHARD_WORKER = r"""
import os, sys, json
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from oracle import pyoracle as O
from suffixarray_b200 import sharded as SH
from suffixarray_b200 import workloads as W

dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
L = 32


def keys_of(text, pos):
    # masked big-endian word of the first 8 bytes (zero after the first terminator / past the end)
    t = np.concatenate([text, np.zeros(8, np.uint8)])
    b = t[pos[:, None] + np.arange(8)[None, :]].astype(np.uint64)
    term = (b == 10) | (b == 255)
    after = (np.cumsum(term, axis=1) - term) > 0
    b[after] = 0
    return (b << (8 * (7 - np.arange(8, dtype=np.uint64)))[None, :]).sum(axis=1).astype(np.uint64)


class CpuStandIn:
    # ShardBuilder's interface with NumPy stages; what the device sort + finish would produce is read
    # off the oracle's array, and every record / halo that arrives through a collective is checked.
    def __init__(self, text, want):
        self.rank, self.world, self.L, self.dev, self.want = rank, world, L, torch.device("cpu"), want
        self.checks = 0
        self.begin(text)

    def begin(self, text, lowercase=False, timing=False, device_text=None):
        self.text, self.n, self.lowercase = text, int(text.size), False
        self.lprime = min(self.L, self.n)
        b = SH.slice_bounds(self.n, self.world)
        self.pos_begin, self.pos_end = b[self.rank], b[self.rank + 1]
        self.m_local = self.pos_end - self.pos_begin
        self.allkeys = keys_of(text, np.arange(self.n))

    def pack(self):
        self.pos_np = np.arange(self.pos_begin, self.pos_end)
        self.keys_np = self.allkeys[self.pos_np]
        return torch.from_numpy(np.bincount((self.keys_np >> np.uint64(48)).astype(np.int64), minlength=SH.NBINS).astype(np.int32))

    def split(self, hist_np):
        self.bin2rank = SH.choose_bin2rank(hist_np, self.world)
        dest = self.bin2rank[(self.keys_np >> np.uint64(48)).astype(np.int64)]
        order = np.argsort(dest, kind="stable")
        self.send_keys = torch.from_numpy(self.keys_np[order].view(np.int64).copy())
        self.send_pos = torch.from_numpy(self.pos_np[order].astype(np.int32))
        self.send_counts = np.bincount(dest, minlength=self.world).astype(np.int64)
        return self.send_counts

    def sort(self, recv_keys, recv_pos, m):
        self.m = int(m)
        pos = recv_pos[:m].numpy().view(np.uint32).astype(np.int64)
        mine = np.flatnonzero(self.bin2rank[(self.allkeys >> np.uint64(48)).astype(np.int64)] == self.rank)
        assert np.array_equal(np.sort(pos), mine), "exchange delivered the wrong records"
        assert np.array_equal(recv_keys[:m].numpy().view(np.uint64), self.allkeys[pos]), "keys and positions came apart"
        # runs arrive in source-rank order, each in ascending position order (the stable split)
        assert np.array_equal(pos, mine), "records are not in (source rank, position) order"
        self.checks += 3

    def _slice(self, all_edges=None):
        counts = np.bincount(self.bin2rank[(self.allkeys >> np.uint64(48)).astype(np.int64)], minlength=self.world)
        off = int(counts[: self.rank].sum())
        return off, self.want[off: off + self.m].astype(np.uint32)

    def edges(self):
        off, sl = self._slice()
        assert np.array_equal(np.sort(sl.astype(np.int64)), np.flatnonzero(
            self.bin2rank[(self.allkeys >> np.uint64(48)).astype(np.int64)] == self.rank)), "a shard is a slice of the array"
        k = min(SH.EDGE, self.m)
        lcp = lambda seg: np.array([0] + [SH.boundary_lcp(self.text, int(a), int(b), self.lprime, False)
                                          for a, b in zip(seg[:-1], seg[1:])], np.uint16)
        hp, tp = sl[:k], sl[self.m - k:]
        tl = lcp(tp)
        if self.m > k:
            tl[0] = SH.boundary_lcp(self.text, int(sl[self.m - k - 1]), int(tp[0]), self.lprime, False)
        return {"m": self.m, "head_pos": hp, "head_lcp": lcp(hp), "tail_pos": tp, "tail_lcp": tl}

    def finish(self, all_edges):
        off, sl = self._slice()
        assert off == sum(int(e["m"]) for e in all_edges[: self.rank])
        if self.m:
            lp, ll, lf, rp, rl = SH.assemble_halos(all_edges, self.rank, self.text, self.lprime, False)
            assert np.array_equal(lp, self.want[off - lp.size: off]), "left halo"
            assert np.array_equal(rp, self.want[off + self.m: off + self.m + rp.size]), "right halo"
            assert lp.size == min(SH.EDGE, off) and rp.size == min(SH.EDGE, self.n - off - self.m)
            if lp.size:
                assert lf == SH.boundary_lcp(self.text, int(self.want[off - 1]), int(self.want[off]), self.lprime, False)
                for i in range(1, lp.size):
                    assert ll[i] == SH.boundary_lcp(self.text, int(lp[i - 1]), int(lp[i]), self.lprime, False), ("left lcp", i)
            for i in range(rp.size):
                prev = int(self.want[off + self.m - 1]) if i == 0 else int(rp[i - 1])
                assert rl[i] == SH.boundary_lcp(self.text, prev, int(rp[i]), self.lprime, False), ("right lcp", i)
            self.checks += 1
        return torch.from_numpy(sl.view(np.int32).copy())

    def adopt_parts(self, gathered, stride, counts):
        g = gathered.numpy().view(np.uint32)
        self.full = np.concatenate([g[i * stride: i * stride + int(c)] for i, c in enumerate(counts)])


ok = True
for text in (W.english_documents(40_000, seed=5), W.company_names_joined(3000, seed=6, upper=False)[0],
             np.full(3000, ord("a"), np.uint8), (np.random.default_rng(3).integers(0, 3, 90) + 97).astype(np.uint8)):
    want = O.oracle_build(text, L)
    b, phases = SH.build_distributed(text, L, 0, builder=CpuStandIn(text, want), exchange="nccl", replicate="allgather")
    ok = ok and np.array_equal(b.full, want) and b.checks >= 3
    flag = torch.tensor([int(ok)])
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ok = bool(flag.item())
if rank == 0:
    print("RESULT" + json.dumps({"ok": ok, "world": world, "phases": sorted(phases)}))
dist.destroy_process_group()
sys.exit(0 if ok else 1)
"""


@pytest.mark.parametrize("world", [2, 3])
def test_gloo_sample_sort_choreography(tmp_path, world):
    """The N > 1 build's host side (suffixarray_b200/sharded.py: build_distributed with the NCCL-style
    exchange) run over gloo on CPU: histogram all-reduce, splitter choice, all-to-all of the (key,
    position) records, edge all-gather + halo assembly, slice all-gather + adopt.  The per-rank device
    stages are replaced by a NumPy stand-in that reads the shard off the oracle's array and checks every
    record and halo that reaches it; the assembled array must be the oracle's."""
    script = tmp_path / "shard_worker.py"
    script.write_text(SHARD_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
                        "--master-addr", "127.0.0.1", "--master-port", str(29750 + world), str(script), ROOT],
                       capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, (r.stdout[-1500:], r.stderr[-3000:])
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("RESULT")][0]
    res = json.loads(line[len("RESULT"):])
    assert res["ok"] and res["world"] == world
    assert {"split", "exchange", "sort+refine", "edges", "finish", "gather", "adopt"} <= set(res["phases"])


def test_array_properties_check_used_above_2_31():
    """bench.py --verify-properties (the checks for indexes too large for any reference array): on the
    oracle's array they hold; a swap of two entries breaks the key order, a duplicate the permutation."""
    import numpy as np
    import torch
    sys.path.insert(0, ROOT)
    import bench
    from oracle import pyoracle as O
    from suffixarray_b200 import workloads as W
    text = W.english_documents(120_000, seed=11)
    want = O.oracle_build(text, 32)
    t = torch.from_numpy(text)
    good = torch.from_numpy(want.view(np.int32).copy())
    perm, order, dg = bench.array_properties(good, t, int(text.size), chunk=1 << 14)
    assert perm and order
    swapped = good.clone()
    swapped[1000], swapped[90_000] = good[90_000].item(), good[1000].item()
    perm2, order2, dg2 = bench.array_properties(swapped, t, int(text.size), chunk=1 << 14)
    assert perm2 and not order2 and not torch.equal(dg, dg2)
    dup = good.clone()
    dup[5] = good[6]
    assert not bench.array_properties(dup, t, int(text.size), chunk=1 << 14)[0]
    raw = torch.from_numpy(np.random.default_rng(1).integers(0, 256, 5000).astype(np.uint8))
    sa_raw = torch.from_numpy(O.oracle_build(raw.numpy(), 32).view(np.int32).copy())
    assert bench.array_properties(sa_raw, raw, 5000)[1] is None          # non-ASCII text: key check not applicable
