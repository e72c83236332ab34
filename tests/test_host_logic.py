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


def test_csv_row_table_of_the_python_mirror(tmp_path):
    """SuffixArray._open_csv (host side of record materialisation) cuts rows at '\\n' outside quotes,
    header = row 0 -- the same definition the device row table uses (csv_rows_kernel): checked here
    against Python's csv module on a file with raw newlines inside quoted cells, ragged rows, CRLF."""
    import csv
    import io
    from suffixarray_b200.suffix_array import SuffixArray
    rows = [["id", "name", "note"], ["1", "plain", "x"], ["2", "two\nlines", 'say "hi"'], ["3", "a,b", "three\n\nlines"],
            ["4"], ["5", "crlf\r\ninside", "y", "extra"], ["6", "last", ""]]
    for term, trailing in (("\n", True), ("\n", False), ("\r\n", True)):
        buf = io.StringIO()
        csv.writer(buf, lineterminator=term).writerows(rows)
        data = buf.getvalue() if trailing else buf.getvalue()[: -len(term)]
        path = tmp_path / "rows.csv"
        path.write_bytes(data.encode())
        sa = SuffixArray.__new__(SuffixArray)                # no index: only the row table is under test
        sa._csv_mm = None
        sa._idx = None
        sa._open_csv(str(path))
        assert sa._row_starts.size == len(rows)
        got = []
        for d in range(len(rows)):
            line = sa._csv_mm[int(sa._row_starts[d]):int(sa._row_ends[d])].decode().rstrip("\r")
            got.append(next(csv.reader([line])))
        assert got == rows
        sa.close()
