"""The native multi-GPU driver (sa_multi.cu: sa_shard_build / sa_build_text_multi): the whole
sample-sort pipeline in one call per rank, small exchanges through a node-local communicator, bulk
data as peer stores.  On a one-GPU box the ranks share the device: as threads of one process
(sa_comm_create_local, raw pointers) and as separate processes (sa_comm_create_shm, CUDA IPC).
Every replica must hold the oracle's array bit for bit and answer queries like a single-GPU index."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import random_text
from oracle import pyoracle as O
from suffixarray_b200 import _lib as S
from suffixarray_b200 import workloads as W

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _cases(rng):
    return [
        (W.english_documents(300_000, seed=41), 32, False),
        (W.company_names_joined(30_000, seed=42, upper=True)[0], 32, True),
        (random_text(rng, 50_000, 2, newline_p=0.02), 16, False),
        (random_text(rng, 20_000, 26), 64, False),
        (rng.integers(0, 256, 30_000).astype(np.uint8), 32, False),       # raw-byte keys (code_bits 8)
        (np.full(5000, ord("a"), np.uint8), 32, False),                   # everything lands on one rank
        (random_text(rng, 700, 3), 8, False),                             # shards smaller than the halo
        (random_text(rng, 40, 2), 32, False),
    ]


@pytest.mark.parametrize("sharded_sa", [False, True])
@pytest.mark.parametrize("world", [2, 3, 8])
def test_threads_of_one_process_match_oracle(world, sharded_sa):
    import torch
    ndev = torch.cuda.device_count()
    rng = np.random.default_rng(60 + world)
    idxs = {}
    for text, L, lower in _cases(rng):
        if L not in idxs:
            idxs[L] = [S.Index(L, device=g % ndev) for g in range(world)]      # handles are reused across builds
        # sharded_sa: no replication -- every handle keeps its key range, queries read the others through peer memory
        S.build_text_multi(idxs[L], text, lowercase=lower, sharded_sa=sharded_sa)
        want_text = O.oracle_lowercase(text) if lower else text
        want = O.oracle_build(want_text, L)
        qs = [bytes(want_text[p:p + 5]).split(b"\n")[0] or b"a" for p in rng.integers(0, max(1, text.size - 5), 20)]
        ws, we = O.oracle_query(want_text, want, L, qs)
        woff, wpos, wdoc = O.oracle_hits(want_text, want, ws, we, 50)
        for g, ix in enumerate(idxs[L]):
            got = ix.sa()
            if not np.array_equal(want, got):
                bad = np.flatnonzero(want != got)
                raise AssertionError(f"world={world} replica {g} n={text.size} L={L}: {bad.size} ranks differ, first {bad[:8]}")
            assert np.array_equal(ix.text(), want_text)
            s, e, off, pos, doc = ix.query(qs, 50)
            assert np.array_equal(s, ws) and np.array_equal(e, we)
            assert np.array_equal(off, woff) and np.array_equal(pos, wpos) and np.array_equal(doc, wdoc)
            assert np.array_equal(ix.doc_starts(), np.concatenate([[0], np.flatnonzero(want_text == 10) + 1]))
    for v in idxs.values():
        for ix in v:
            ix.close()


WORKER = r"""
import json, os, sys
import numpy as np
sys.path.insert(0, sys.argv[1])
rank, world, name = int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
import torch
from oracle import pyoracle as O
from suffixarray_b200 import _lib as S
from suffixarray_b200 import workloads as W
dev = rank % torch.cuda.device_count()
comm = S.Comm(name, rank, world)
idx = S.Index(32, device=dev)
ok = True
for rep, (n_chars, seed, lower) in enumerate([(400_000, 5, False), (150_000, 6, True), (900_000, 7, False)]):
    text = W.english_documents(n_chars, seed=seed) if not lower else W.company_names_joined(9000, seed=seed, upper=True)[0]
    n = int(text.size)
    a, b = rank * n // world, (rank + 1) * n // world             # this rank uploads ONLY its slice
    mine = np.ascontiguousarray(text[a:b])
    # the second build keeps the array distributed (SA_FLAG_SHARDED_SA): queries and sa() read the peers' shards
    flags = (S.SA_FLAG_LOWERCASE if lower else 0) | (S.SA_FLAG_SHARDED_SA if rep == 1 else 0)
    ph = S.shard_build(idx, comm, mine.ctypes.data, a, b - a, n, flags)
    want_text = O.oracle_lowercase(text) if lower else text
    want = O.oracle_build(want_text, 32)
    ok = ok and bool(np.array_equal(idx.sa(), want)) and bool(np.array_equal(idx.text(), want_text))
    qb, qo = W.substring_queries(want_text, 2000, seed=9)
    qs = W.unpack_queries(qb, qo)
    ws, we = O.oracle_query(want_text, want, 32, qs)
    s, e, _ = idx.query_ranges(qs)
    ok = ok and bool(np.array_equal(s, ws) and np.array_equal(e, we))
    comm.barrier()          # nobody rebuilds while a peer still reads this build's shard
comm.barrier()
idx.close()
comm.close()
print("RESULT" + json.dumps({"rank": rank, "ok": ok, "phases": ph}))
"""


@pytest.mark.parametrize("world", [2, 4])
def test_processes_over_shared_memory_and_ipc(world, tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    name = f"/sa_test_{os.getpid()}_{world}"
    procs = [subprocess.Popen([sys.executable, str(script), ROOT, str(r), str(world), name], stdout=subprocess.PIPE,
                              stderr=subprocess.PIPE, text=True) for r in range(world)]
    outs = []
    for p in procs:
        try:
            o, e = p.communicate(timeout=600)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        assert p.returncode == 0, e[-3000:]
        line = [ln for ln in o.splitlines() if ln.startswith("RESULT")][-1]
        outs.append(json.loads(line[len("RESULT"):]))
    assert all(o["ok"] for o in outs), outs


@pytest.mark.parametrize("flavour", ["ctypes", "cython"])
def test_class_devices_keyword(flavour, tmp_path):
    """SuffixArray(documents=..., devices=[0, 1, ...]): one index over the GPUs of this process, no
    torch.distributed in the caller; query batches are split over the replicas."""
    import torch
    ndev = torch.cuda.device_count()
    if ndev < 2:
        pytest.skip("needs 2 GPUs in one process")
    if flavour == "cython":
        from suffixarray_b200.suffix_array_cy import SuffixArray
    else:
        from suffixarray_b200 import SuffixArray
    names = W.company_names(60_000, seed=77, upper=True)
    docs = [b.decode() for b in names]
    devs = list(range(min(ndev, 4)))
    multi = SuffixArray(documents=docs, max_suffix_length=32, devices=devs)
    single = SuffixArray(documents=docs, max_suffix_length=32, device=0)
    assert np.array_equal(multi.suffix_array, single.suffix_array)
    qs = ["inc", "group", docs[5].lower(), docs[17][:4].lower(), "qqqq", "", "llc", "a", "holdings", "co", "corp", "sa"]
    assert multi.query_records_batch(qs, k=25) == single.query_records_batch(qs, k=25)
    assert multi.query_records("inc", k=5) == single.query_records("inc", k=5)
    data = W.csv_file_bytes(3000, seed=3, quoted_fraction=0.1)
    p = str(tmp_path / "d.csv")
    with open(p, "wb") as f:
        f.write(data)
    a = SuffixArray(csv_file=p, search_column="name", max_suffix_length=32, devices=devs[:2])
    b = SuffixArray(csv_file=p, search_column="name", max_suffix_length=32)
    qs2 = ["inc", "llc", "corp", "a", "group", "zz"]
    assert a.query_records_batch(qs2, k=10) == b.query_records_batch(qs2, k=10)
    for x in (multi, single, a, b):
        x.close()
