"""ctypes bindings for the parity checker.  TEST INFRASTRUCTURE ONLY.

Two libraries, both built by ``oracle/Makefile``:

* ``oracle/libsa_oracle.so`` -- our plain-C restatement (``sa_oracle.c``) of the reference's
  ``construct_truncated_suffix_array`` (engine.c:837-866, :696-812, :664-677),
  ``get_substring_positions`` (engine.c:869-918), the hit expansion (engine.c:1185-1187)
  and the CSV column scan (engine.c:522-605, :632-651).
* ``oracle/_ref/libsa_ref.so`` -- the UNMODIFIED reference sources compiled from
  ``/root/reference`` (only buildable in the authoring container; the built file travels to
  the GPU box).  ``have_ref()`` says whether it is present.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs import this
module.  Nothing under ``suffixarray_b200/`` does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_ORACLE_SO = os.path.join(_HERE, "libsa_oracle.so")
_REF_SO = os.path.join(_HERE, "_ref", "libsa_ref.so")
REF_PAD = 512

_u8p = C.POINTER(C.c_uint8)
_u32p = C.POINTER(C.c_uint32)
_u64p = C.POINTER(C.c_uint64)
_i32p = C.POINTER(C.c_int32)


def build(force: bool = False) -> None:
    """Compile the checker (and oracle/_ref when /root/reference is present)."""
    if force or not os.path.exists(_ORACLE_SO) or (
        os.path.exists("/root/reference/suffix_array/engine.c") and not os.path.exists(_REF_SO)
    ):
        subprocess.run(["make", "-C", _HERE, "all"], check=True, capture_output=True)


def _p(a: np.ndarray, t):
    return a.ctypes.data_as(t)


_oracle = None
_ref = None


def oracle_lib():
    global _oracle
    if _oracle is None:
        build()
        lib = C.CDLL(_ORACLE_SO)
        lib.sao_build.argtypes = [_u8p, C.c_uint64, C.c_uint32, _u32p]
        lib.sao_build.restype = C.c_int
        lib.sao_query_batch.argtypes = [_u8p, C.c_uint64, _u32p, C.c_uint32, _u8p, _u64p, C.c_uint64,
                                        _u32p, _u32p, C.c_int]
        lib.sao_query_batch.restype = C.c_int
        lib.sao_query_batch_padded.argtypes = lib.sao_query_batch.argtypes
        lib.sao_query_batch_padded.restype = C.c_int
        lib.sao_hit_offsets.argtypes = [_u32p, _u32p, C.c_uint64, C.c_uint32, _u64p]
        lib.sao_hit_offsets.restype = None
        lib.sao_hits.argtypes = [_u8p, C.c_uint64, _u32p, _u32p, _u32p, C.c_uint64, C.c_uint32, _u64p,
                                 _u32p, _u32p]
        lib.sao_hits.restype = C.c_int
        lib.sao_lowercase.argtypes = [_u8p, C.c_uint64]
        lib.sao_lowercase.restype = None
        lib.sao_csv_scan.argtypes = [_u8p, C.c_uint64, C.c_uint32, C.c_uint32, _u8p, _u32p, _u8p]
        lib.sao_csv_scan.restype = C.c_uint64
        lib.sao_csv_remap.argtypes = [_u32p, C.c_uint64, _u32p, _u8p, _u8p]
        lib.sao_csv_remap.restype = None
        lib.sao_build_tail.argtypes = [_u8p, C.c_uint64, C.c_uint64, C.c_uint32, _u32p]
        lib.sao_build_tail.restype = C.c_int
        lib.sao_query_batch_tail.argtypes = [_u8p, C.c_uint64, C.c_uint64, _u32p, C.c_uint32, _u8p, _u64p, C.c_uint64,
                                             _u32p, _u32p]
        lib.sao_query_batch_tail.restype = C.c_int
        lib.sao_query_file_batch.argtypes = [_u8p, C.c_uint64, _u32p, C.c_uint64, C.c_uint32, _u8p, _u64p, C.c_uint64,
                                             _u32p, _u32p, C.c_int]
        lib.sao_query_file_batch.restype = C.c_int
        lib.sao_csv_row_starts.argtypes = [_u8p, C.c_uint64, _u64p]
        lib.sao_csv_row_starts.restype = C.c_uint64
        lib.sao_dedupe.argtypes = [_u64p, C.c_uint64, _u32p, _u32p, _u64p, _u32p, _u32p]
        lib.sao_dedupe.restype = C.c_int
        _oracle = lib
    return _oracle


def have_ref() -> bool:
    if not os.path.exists(_REF_SO) and os.path.exists("/root/reference/suffix_array/engine.c"):
        build()
    return os.path.exists(_REF_SO)


def ref_lib():
    global _ref
    if _ref is None:
        if not have_ref():
            raise RuntimeError("oracle/_ref/libsa_ref.so is not built (needs /root/reference)")
        lib = C.CDLL(_REF_SO)
        lib.ref_build.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, _u32p]
        lib.ref_build.restype = None
        lib.ref_query.argtypes = [C.c_char_p, _u32p, C.c_uint32, C.c_uint32, C.c_char_p, _u32p, _u32p]
        lib.ref_query.restype = C.c_int
        lib.ref_query_batch.argtypes = [C.c_void_p, _u32p, C.c_uint32, C.c_uint32, C.c_void_p, _u64p,
                                        C.c_uint64, _u32p, _u32p, _u8p, C.c_int]
        lib.ref_query_batch.restype = None
        lib.ref_query_records_batch.argtypes = [C.c_void_p, _u32p, C.c_uint32, C.c_uint32, C.c_void_p, _u64p, C.c_uint64,
                                                C.c_uint32, _u32p, C.c_int]
        lib.ref_query_records_batch.restype = C.c_uint64
        lib.ref_csv_build.argtypes = [C.c_char_p, C.c_char_p, C.c_uint32, _u32p, _u8p, C.c_uint64]
        lib.ref_csv_build.restype = C.c_uint32
        lib.ref_libsais.argtypes = [_u8p, _i32p, C.c_int32, C.c_int]
        lib.ref_libsais.restype = C.c_int
        lib.ref_libsais64.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int]
        lib.ref_libsais64.restype = C.c_int64
        lib.ref_max_threads.argtypes = []
        lib.ref_max_threads.restype = C.c_int
        lib.ref_set_threads.argtypes = [C.c_int]
        lib.ref_set_threads.restype = None
        lib.ref_csv_open.argtypes = [C.c_char_p, C.c_char_p, C.c_uint32]
        lib.ref_csv_open.restype = C.c_void_p
        lib.ref_csv_n.argtypes = [C.c_void_p]
        lib.ref_csv_n.restype = C.c_uint32
        lib.ref_csv_partitions.argtypes = [C.c_void_p]
        lib.ref_csv_partitions.restype = C.c_uint32
        lib.ref_csv_part_n.argtypes = [C.c_void_p, C.c_uint32]
        lib.ref_csv_part_n.restype = C.c_uint32
        lib.ref_csv_part_bounds.argtypes = [C.c_void_p, C.c_uint32, _u64p, _u64p]
        lib.ref_csv_part_bounds.restype = None
        lib.ref_csv_part_copy.argtypes = [C.c_void_p, C.c_uint32, _u32p, _u8p]
        lib.ref_csv_part_copy.restype = None
        lib.ref_csv_copy.argtypes = [C.c_void_p, _u32p, _u8p]
        lib.ref_csv_copy.restype = None
        lib.ref_csv_close.argtypes = [C.c_void_p]
        lib.ref_csv_close.restype = None
        lib.ref_csv_query_batch.argtypes = [C.c_void_p, C.c_char_p, C.c_void_p, _u64p, C.c_uint64, _u32p, _u32p, _u8p, C.c_int]
        lib.ref_csv_query_batch.restype = None
        lib.ref_csv_records.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_uint32, C.c_void_p, C.c_uint64, _u64p]
        lib.ref_csv_records.restype = C.c_int64
        _ref = lib
    return _ref


def _as_u8(text) -> np.ndarray:
    if isinstance(text, (bytes, bytearray, memoryview)):
        return np.frombuffer(bytes(text), dtype=np.uint8)
    return np.ascontiguousarray(text, dtype=np.uint8)


def pack_queries(queries):
    """list[bytes] -> (qbytes u8, qoff u64[nq+1])."""
    lens = np.fromiter((len(q) for q in queries), dtype=np.uint64, count=len(queries))
    qoff = np.zeros(len(queries) + 1, dtype=np.uint64)
    np.cumsum(lens, out=qoff[1:])
    qbytes = np.frombuffer(b"".join(queries), dtype=np.uint8).copy() if len(queries) else np.zeros(0, np.uint8)
    if qbytes.size == 0:
        qbytes = np.zeros(1, np.uint8)
    return qbytes, qoff


# ----------------------------------------------------------------------------- oracle (port)
def oracle_build(text, L: int) -> np.ndarray:
    t = _as_u8(text)
    sa = np.empty(t.size, dtype=np.uint32)
    tt = t if t.size else np.zeros(1, np.uint8)
    rc = oracle_lib().sao_build(_p(tt, _u8p), t.size, L, _p(sa if sa.size else np.zeros(1, np.uint32), _u32p))
    if rc != 0:
        raise RuntimeError(f"sao_build failed: {rc}")
    return sa


def oracle_build_tail(text, n: int, L: int) -> np.ndarray:
    """ORDER-E of the first n positions of `text`; the bytes after n are look-ahead (a partition of a
    longer text, suffix_array.pyx:176-180)."""
    t = _as_u8(text)
    sa = np.empty(max(n, 1), dtype=np.uint32)
    rc = oracle_lib().sao_build_tail(_p(t, _u8p), n, t.size - n, L, _p(sa, _u32p))
    if rc != 0:
        raise RuntimeError(f"sao_build_tail failed: {rc}")
    return sa[:n]


def oracle_query_tail(text, n: int, sa: np.ndarray, L: int, queries):
    t = _as_u8(text)
    qbytes, qoff = pack_queries(queries)
    nq = len(queries)
    start = np.empty(max(nq, 1), dtype=np.uint32)
    end = np.empty(max(nq, 1), dtype=np.uint32)
    rc = oracle_lib().sao_query_batch_tail(_p(t, _u8p), n, t.size - n, _p(np.ascontiguousarray(sa, np.uint32), _u32p), L,
                                           _p(qbytes, _u8p), _p(qoff, _u64p), nq, _p(start, _u32p), _p(end, _u32p))
    if rc != 0:
        raise RuntimeError("sao_query_batch_tail failed")
    return start[:nq], end[:nq]


def oracle_query(text, sa: np.ndarray, L: int, queries, nthreads: int = 0):
    """-> (start u32[nq], end u32[nq]) with the reference's UINT32_MAX sentinels."""
    t = _as_u8(text)
    qbytes, qoff = pack_queries(queries)
    nq = len(queries)
    start = np.empty(max(nq, 1), dtype=np.uint32)
    end = np.empty(max(nq, 1), dtype=np.uint32)
    tt = t if t.size else np.zeros(1, np.uint8)
    sa_ = np.ascontiguousarray(sa, dtype=np.uint32) if sa.size else np.zeros(1, np.uint32)
    rc = oracle_lib().sao_query_batch(_p(tt, _u8p), t.size, _p(sa_, _u32p), L, _p(qbytes, _u8p),
                                      _p(qoff, _u64p), nq, _p(start, _u32p), _p(end, _u32p), nthreads)
    if rc != 0:
        raise RuntimeError("sao_query_batch failed")
    return start[:nq], end[:nq]


def oracle_hits(text, sa: np.ndarray, start: np.ndarray, end: np.ndarray, k: int):
    """-> (hit_off u64[nq+1], hit_pos u32[total], hit_doc u32[total])."""
    t = _as_u8(text)
    nq = start.size
    start = np.ascontiguousarray(start, dtype=np.uint32)
    end = np.ascontiguousarray(end, dtype=np.uint32)
    hit_off = np.zeros(nq + 1, dtype=np.uint64)
    s_ = start if nq else np.zeros(1, np.uint32)
    e_ = end if nq else np.zeros(1, np.uint32)
    oracle_lib().sao_hit_offsets(_p(s_, _u32p), _p(e_, _u32p), nq, k, _p(hit_off, _u64p))
    total = int(hit_off[nq])
    hit_pos = np.empty(max(total, 1), dtype=np.uint32)
    hit_doc = np.empty(max(total, 1), dtype=np.uint32)
    tt = t if t.size else np.zeros(1, np.uint8)
    sa_ = np.ascontiguousarray(sa, dtype=np.uint32) if sa.size else np.zeros(1, np.uint32)
    rc = oracle_lib().sao_hits(_p(tt, _u8p), t.size, _p(sa_, _u32p), _p(s_, _u32p), _p(e_, _u32p), nq, k,
                               _p(hit_off, _u64p), _p(hit_pos, _u32p), _p(hit_doc, _u32p))
    if rc != 0:
        raise RuntimeError("sao_hits failed")
    return hit_off, hit_pos[:total], hit_doc[:total]


def oracle_lowercase(text) -> np.ndarray:
    t = _as_u8(text).copy()
    if t.size:
        oracle_lib().sao_lowercase(_p(t, _u8p), t.size)
    return t


def oracle_csv_scan(file_bytes: bytes, column_idx: int, num_columns: int):
    """-> (text u8[n], map u32[n], quoted u8[n]) for a single-partition file."""
    size = len(file_bytes)
    f = np.zeros(size + 8, dtype=np.uint8)
    f[:size] = np.frombuffer(file_bytes, dtype=np.uint8)
    text = np.zeros(size + 2, dtype=np.uint8)
    mp = np.zeros(size + 2, dtype=np.uint32)
    qd = np.zeros(size + 2, dtype=np.uint8)
    n = oracle_lib().sao_csv_scan(_p(f, _u8p), size, column_idx, num_columns, _p(text, _u8p), _p(mp, _u32p),
                                  _p(qd, _u8p))
    return text[:n].copy(), mp[:n].copy(), qd[:n].copy()


def oracle_csv_remap(sa: np.ndarray, mp: np.ndarray, quoted_by_text: np.ndarray):
    sa2 = np.ascontiguousarray(sa, dtype=np.uint32).copy()
    qr = np.zeros(max(sa2.size, 1), dtype=np.uint8)
    if sa2.size:
        oracle_lib().sao_csv_remap(_p(sa2, _u32p), sa2.size, _p(np.ascontiguousarray(mp, np.uint32), _u32p),
                                   _p(np.ascontiguousarray(quoted_by_text, np.uint8), _u8p), _p(qr, _u8p))
    return sa2, qr[: sa2.size]


def oracle_query_file(file_bytes: bytes, sa_file: np.ndarray, L: int, queries, nthreads: int = 0):
    """get_substring_positions_file (engine.c:920-999) restated: probes compare raw file bytes."""
    size = len(file_bytes)
    f = np.zeros(size + 8, dtype=np.uint8)
    f[:size] = np.frombuffer(file_bytes, dtype=np.uint8)
    qbytes, qoff = pack_queries(queries)
    nq = len(queries)
    start = np.empty(max(nq, 1), dtype=np.uint32)
    end = np.empty(max(nq, 1), dtype=np.uint32)
    sa_ = np.ascontiguousarray(sa_file, dtype=np.uint32) if sa_file.size else np.zeros(1, np.uint32)
    rc = oracle_lib().sao_query_file_batch(_p(f, _u8p), size, _p(sa_, _u32p), sa_file.size, L, _p(qbytes, _u8p),
                                           _p(qoff, _u64p), nq, _p(start, _u32p), _p(end, _u32p), nthreads)
    if rc != 0:
        raise RuntimeError("sao_query_file_batch failed")
    return start[:nq], end[:nq]


def oracle_csv_row_starts(file_bytes: bytes) -> np.ndarray:
    """-> u64[rows + 1]: first byte of every row of the file (row 0 = header), then the file size."""
    size = len(file_bytes)
    f = np.zeros(size + 8, dtype=np.uint8)
    f[:size] = np.frombuffer(file_bytes, dtype=np.uint8)
    out = np.zeros(size + 2, dtype=np.uint64)
    rows = oracle_lib().sao_csv_row_starts(_p(f, _u8p), size, _p(out, _u64p))
    return out[: rows + 1].copy()


def oracle_hits_rows(sa_file: np.ndarray, row_starts: np.ndarray, start: np.ndarray, end: np.ndarray, k: int):
    """CSV hit expansion: hit_pos = file offset SA[rank], hit_doc = row of the file holding it."""
    nq = start.size
    hit_off = np.zeros(nq + 1, dtype=np.uint64)
    s_ = np.ascontiguousarray(start, np.uint32) if nq else np.zeros(1, np.uint32)
    e_ = np.ascontiguousarray(end, np.uint32) if nq else np.zeros(1, np.uint32)
    oracle_lib().sao_hit_offsets(_p(s_, _u32p), _p(e_, _u32p), nq, k, _p(hit_off, _u64p))
    pos = [sa_file[int(start[i]): int(start[i]) + int(hit_off[i + 1] - hit_off[i])] for i in range(nq)]
    hit_pos = np.concatenate(pos).astype(np.uint32) if pos else np.zeros(0, np.uint32)
    hit_doc = (np.searchsorted(row_starts[:-1], hit_pos.astype(np.uint64), side="right") - 1).astype(np.uint32)
    return hit_off, hit_pos, hit_doc


def oracle_dedupe(hit_off: np.ndarray, hit_pos: np.ndarray, hit_doc: np.ndarray):
    """-> (off, pos, doc): per query the first hit (rank order) of every record id."""
    nq = hit_off.size - 1
    total = int(hit_off[-1])
    off = np.zeros(nq + 1, dtype=np.uint64)
    pos = np.zeros(max(total, 1), dtype=np.uint32)
    doc = np.zeros(max(total, 1), dtype=np.uint32)
    hp = np.ascontiguousarray(hit_pos, np.uint32) if total else np.zeros(1, np.uint32)
    hd = np.ascontiguousarray(hit_doc, np.uint32) if total else np.zeros(1, np.uint32)
    rc = oracle_lib().sao_dedupe(_p(np.ascontiguousarray(hit_off, np.uint64), _u64p), nq, _p(hp, _u32p), _p(hd, _u32p),
                                 _p(off, _u64p), _p(pos, _u32p), _p(doc, _u32p))
    if rc != 0:
        raise RuntimeError("sao_dedupe failed")
    u = int(off[-1])
    return off, pos[:u], doc[:u]


# ----------------------------------------------------------------------------- reference
class RefCsvIndex:
    """The reference's own CSV index (engine.c:1454) kept alive for its query functions."""

    def __init__(self, path: str, column: str, L: int):
        self.path = path
        self.h = ref_lib().ref_csv_open(path.encode(), column.encode(), L)
        if not self.h:
            raise RuntimeError("ref_csv_open failed")
        self.n = int(ref_lib().ref_csv_n(self.h))

    def arrays(self):
        sa = np.zeros(max(self.n, 1), dtype=np.uint32)
        qd = np.zeros(max(self.n, 1), dtype=np.uint8)
        ref_lib().ref_csv_copy(self.h, _p(sa, _u32p), _p(qd, _u8p))
        return sa[: self.n], qd[: self.n]

    def partitions(self) -> int:
        return int(ref_lib().ref_csv_partitions(self.h))

    def partition(self, p: int):
        """-> (start, end, sa u32, quoted u8) of partition p (files >= 2 GiB have several)."""
        n = int(ref_lib().ref_csv_part_n(self.h, p))
        a, b = np.zeros(1, np.uint64), np.zeros(1, np.uint64)
        ref_lib().ref_csv_part_bounds(self.h, p, _p(a, _u64p), _p(b, _u64p))
        sa = np.zeros(max(n, 1), dtype=np.uint32)
        qd = np.zeros(max(n, 1), dtype=np.uint8)
        ref_lib().ref_csv_part_copy(self.h, p, _p(sa, _u32p), _p(qd, _u8p))
        return int(a[0]), int(b[0]), sa[:n], qd[:n]

    def query(self, queries, nthreads: int = 0):
        """get_substring_positions_file -> (start, end, skipped)."""
        nq = len(queries)
        qbytes, qoff = pack_queries([q + b"\0" for q in queries])
        start = np.full(max(nq, 1), 0xFFFFFFFF, dtype=np.uint32)
        end = np.full(max(nq, 1), 0xFFFFFFFF, dtype=np.uint32)
        skipped = np.zeros(max(nq, 1), dtype=np.uint8)
        ref_lib().ref_csv_query_batch(self.h, self.path.encode(), qbytes.ctypes.data, _p(qoff, _u64p), nq,
                                      _p(start, _u32p), _p(end, _u32p), _p(skipped, _u8p), nthreads)
        return start[:nq], end[:nq], skipped[:nq]

    def records(self, query: bytes, k: int, cap: int = 1 << 22):
        """get_matching_records_file -> list[bytes], or None when the shim fenced the query."""
        out = np.zeros(cap, dtype=np.uint8)
        off = np.zeros(k + 2, dtype=np.uint64)
        got = ref_lib().ref_csv_records(self.h, self.path.encode(), query, k, out.ctypes.data, cap, _p(off, _u64p))
        if got < 0:
            return None
        b = out.tobytes()
        return [b[int(off[j]): int(off[j + 1])] for j in range(int(got))]

    def close(self):
        if self.h:
            ref_lib().ref_csv_close(self.h)
            self.h = None


def _padded(text) -> np.ndarray:
    t = _as_u8(text)
    out = np.zeros(t.size + REF_PAD, dtype=np.uint8)
    out[: t.size] = t
    return out


def ref_build(text, L: int) -> np.ndarray:
    t = _as_u8(text)
    tp = _padded(t)
    sa = np.empty(max(t.size, 1), dtype=np.uint32)
    if t.size:
        ref_lib().ref_build(tp.ctypes.data_as(C.c_char_p), t.size, L, _p(sa, _u32p))
    return sa[: t.size]


def ref_build_padded(text_padded: np.ndarray, n: int, L: int, sa_out: np.ndarray) -> None:
    """Timed-baseline entry: no copies; text_padded must hold >= REF_PAD zero bytes after n."""
    ref_lib().ref_build(text_padded.ctypes.data_as(C.c_char_p), n, L, _p(sa_out, _u32p))


def ref_query(text, sa: np.ndarray, L: int, queries, nthreads: int = 0, text_is_padded_n: int | None = None):
    """-> (start, end, skipped).  Queries must not contain NUL (they are passed as C strings).
    skipped[i] = 1 where the reference would take its out-of-bounds wrap (see ref_shim.c)."""
    if text_is_padded_n is None:
        t = _as_u8(text)
        n = t.size
        tp = _padded(t)
    else:
        tp = text
        n = text_is_padded_n
    nq = len(queries)
    zq = [q + b"\0" for q in queries]
    qbytes, qoff = pack_queries(zq)
    start = np.full(max(nq, 1), 0xFFFFFFFF, dtype=np.uint32)
    end = np.full(max(nq, 1), 0xFFFFFFFF, dtype=np.uint32)
    skipped = np.zeros(max(nq, 1), dtype=np.uint8)
    sa_ = np.ascontiguousarray(sa, dtype=np.uint32)
    ref_lib().ref_query_batch(tp.ctypes.data, _p(sa_, _u32p), n, L, qbytes.ctypes.data, _p(qoff, _u64p), nq,
                              _p(start, _u32p), _p(end, _u32p), _p(skipped, _u8p), nthreads)
    return start[:nq], end[:nq], skipped[:nq]


def ref_csv_build(path: str, column: str, L: int, cap: int):
    sa = np.zeros(cap, dtype=np.uint32)
    qd = np.zeros(cap, dtype=np.uint8)
    n = ref_lib().ref_csv_build(path.encode(), column.encode(), L, _p(sa, _u32p), _p(qd, _u8p), cap)
    if n > cap:
        raise RuntimeError(f"cap too small: {n} > {cap}")
    return sa[:n].copy(), qd[:n].copy()


def ref_libsais(text, threads: int = 1) -> np.ndarray:
    t = _as_u8(text)
    sa = np.empty(max(t.size, 1), dtype=np.int32)
    rc = ref_lib().ref_libsais(_p(t, _u8p), _p(sa, _i32p), t.size, threads)
    if rc != 0:
        raise RuntimeError(f"libsais failed: {rc}")
    return sa[: t.size]


def ref_libsais64(text, threads: int = 0) -> np.ndarray:
    """libsais64[_omp] of the reference tree: int64 FULL suffix array (timing baseline B at n >= 2^31)."""
    t = _as_u8(text)
    sa = np.empty(max(t.size, 1), dtype=np.int64)
    rc = ref_lib().ref_libsais64(t.ctypes.data, sa.ctypes.data, t.size, threads)
    if rc != 0:
        raise RuntimeError(f"libsais64 failed: {rc}")
    return sa[: t.size]


def ref_max_threads() -> int:
    return int(ref_lib().ref_max_threads())


def ref_use_all_cores() -> int:
    """Set the OpenMP thread count of the reference library to the cores this process may run on
    (torchrun exports OMP_NUM_THREADS=1 to its workers) -> that count."""
    import os
    n = len(os.sched_getaffinity(0))
    ref_lib().ref_set_threads(n)
    return n
