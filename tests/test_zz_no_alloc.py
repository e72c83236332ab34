"""sa_hit_offsets_noalloc -- the entry point that allocates nothing (include/sa_cuda.h; no counterpart in
the reference): its scratch is a fixed __device__ array
inside the library.  Checked through the C-ABI against the existing entry point (sa_query_ranges_r, whose
hit_off comes from the cudaMalloc-backed scan of query.cu) on calls in a row that share the scratch --
integer work, so the outputs must be identical.

Collected last on purpose: passes alone in a fresh process and after
the other tests."""
import numpy as np
import pytest

from suffixarray_b200 import _lib as S
from suffixarray_b200 import workloads as W

pytestmark = pytest.mark.gpu


def _index_and_batch(nq, seed):
    text = W.english_documents(3_000_000, seed=seed)
    qbytes, qoff = W.substring_queries(text, nq, seed=seed + 1)
    idx = S.Index(32)
    idx.build_text(text)
    return idx, qbytes, qoff


def _noalloc(idx, torch, s, e, k, stream=0):
    nq = int(s.size)
    d_s = torch.from_numpy(s.view(np.int32)).cuda()
    d_e = torch.from_numpy(e.view(np.int32)).cuda()
    d_off = torch.full((nq + 1,), -1, dtype=torch.int64, device="cuda")       # stale contents must all be overwritten
    torch.cuda.synchronize()
    idx.hit_offsets_noalloc(d_s.data_ptr(), d_e.data_ptr(), nq, k, d_off.data_ptr(), stream)
    torch.cuda.synchronize()
    return d_off.cpu().numpy().view(np.uint64)


def test_noalloc_hit_offsets_equal_existing_entry_point():
    import torch
    torch.cuda.set_device(0)
    idx, qbytes, qoff = _index_and_batch(200_000, seed=9100)
    res = idx.result()
    # call 1: the whole batch, k = 100 (196 block sums + total in the scratch)
    s, e, off = _ranges(res, qbytes, qoff, 100)
    assert int(off[-1]) > s.size                                   # long lists exist, k clamps some of them
    got = _noalloc(idx, torch, s, e, 100)
    assert np.array_equal(got, off.astype(np.uint64))
    # call 2, straight after: a shorter batch with another k -- the scratch still holds call 1's sums
    m = 70_001
    s2, e2, off2 = _ranges(res, qbytes, qoff[: m + 1], 7)
    assert np.array_equal(s2, s[:m]) and np.array_equal(e2, e[:m])
    got2 = _noalloc(idx, torch, s2, e2, 7)
    assert np.array_equal(got2, off2.astype(np.uint64))
    assert int(got2[-1]) < int(got[m])
    # call 3: the first batch again, on a non-default stream
    st = torch.cuda.Stream()
    got3 = _noalloc(idx, torch, s, e, 100, stream=st.cuda_stream)
    assert np.array_equal(got3, off.astype(np.uint64))
    # misses only, a single query, an empty batch
    miss = np.full(5, 0xFFFFFFFF, dtype=np.uint32)
    assert np.array_equal(_noalloc(idx, torch, miss, miss, 10), np.zeros(6, dtype=np.uint64))
    one = min(3, int(e[0]) - int(s[0]) + 1) if s[0] != 0xFFFFFFFF else 0
    assert np.array_equal(_noalloc(idx, torch, s[:1].copy(), e[:1].copy(), 3), np.array([0, one], dtype=np.uint64))
    assert np.array_equal(_noalloc(idx, torch, s[:0].copy(), e[:0].copy(), 3), np.zeros(1, dtype=np.uint64))
    res.close()
    idx.close()


def test_noalloc_rejects_batches_beyond_the_scratch():
    import torch
    torch.cuda.set_device(0)
    idx = S.Index(32)
    idx.build_text(b"\nabc\nabd")
    d = torch.zeros(8, dtype=torch.int64, device="cuda")
    with pytest.raises(S.SaCudaError) as ei:
        idx.hit_offsets_noalloc(d.data_ptr(), d.data_ptr(), S.SA_NOALLOC_MAX_QUERIES + 1, 1, d.data_ptr())
    assert ei.value.code == -5 and "SA_NOALLOC_MAX_QUERIES" in str(ei.value)
    with pytest.raises(S.SaCudaError):
        idx.hit_offsets_noalloc(d.data_ptr(), d.data_ptr(), 4, 1, 0)          # NULL output
    unbuilt = S.Index(32)
    with pytest.raises(S.SaCudaError):
        unbuilt.hit_offsets_noalloc(d.data_ptr(), d.data_ptr(), 4, 1, d.data_ptr())
    unbuilt.close()
    idx.close()


def _ranges(res, qbytes, qoff, k):
    nq = int(qoff.size - 1)
    qb = np.ascontiguousarray(qbytes[: int(qoff[nq])])
    return res.ranges_packed(qb, np.ascontiguousarray(qoff), k=k)
