# cython: language_level=3
"""The reference's Cython ``SuffixArray`` class (suffix_array/suffix_array.pyx:110-267), re-pointed at
libsa_cuda.so: the ``cdef extern from "engine.h"`` block (pyx:37-101) becomes
``cdef extern from "sa_cuda.h"``, ``engine.c`` is no longer compiled into the extension, and the
class keeps its names, keyword arguments and error behaviour:

* ``SuffixArray(max_suffix_length=64)``                                  pyx:125
* ``SuffixArray(documents=docs, max_suffix_length=32)``                  README.md:23
* ``SuffixArray(csv_file=f, search_column=c, max_suffix_length=32)``     README.md:40-44
* ``construct_truncated_suffix_array_documents(documents)``              pyx:129-180
* ``construct_truncated_suffix_array_from_csv(filename, search_column)`` pyx:183-207
  (``search_column``: one name or a list -- pyx:197 "TODO: Allow multiple columns")
* ``query_records(substring, k=1000)``                                   pyx:209-267
* ``query_records_batch(substrings, k=1000)``   the batched entry the GPU path wants

Builds and queries run with the GIL released (as pyx:159,175,179,200 do for the builds).  All index
arithmetic is in the CUDA library; this file joins / encodes strings and cuts records.  No CPU
fallback: without a CUDA device every constructor call raises ``SaCudaError``.
The class logic itself lives once in ``suffixarray_b200/_host.py`` (shared with the ctypes class in
``suffix_array.py``); this module is the compiled binding: the ``cdef extern`` block, the handle type
and the GIL-free calls.  Both classes are run through the same drop-in tests.  Built by ``__graft_entry__.build()`` (cython + gcc, linked
against ``libsa_cuda.so`` with an ``$ORIGIN`` rpath).
"""
from libc.stdint cimport uint8_t, uint32_t, uint64_t, uintptr_t

import numpy as np

from suffixarray_b200._host import SuffixArrayHost

cdef extern from "sa_cuda.h" nogil:
    ctypedef struct sa_index:
        pass
    const char* sa_last_error()
    int sa_create(sa_index** out, uint32_t max_suffix_length, int device)
    void sa_destroy(sa_index* idx)
    ctypedef struct sa_result:
        pass
    int sa_build_text_tail(sa_index* idx, const uint8_t* text, uint64_t n, uint64_t tail, uint32_t flags)
    int sa_build_csv(sa_index* idx, const char* csv_path, const char* column, uint32_t flags)
    int sa_build_csv_partition(sa_index* idx, const char* csv_path, const char* column, uint64_t partition_start,
                               uint32_t flags, uint64_t* partition_end, uint64_t* file_bytes)
    int sa_build_csv_columns(sa_index** idxs, const char** columns, uint32_t ncols, const char* csv_path, uint32_t flags)
    uint64_t sa_size(const sa_index* idx)
    int sa_copy_sa_u32(sa_index* idx, uint32_t* out, uint64_t cap)
    int sa_copy_text(sa_index* idx, uint8_t* out, uint64_t cap)
    int sa_result_create(sa_index* idx, sa_result** out)
    void sa_result_destroy(sa_result* res)
    int sa_query_ranges_r(sa_index* idx, sa_result* res, const uint8_t* qbytes, const uint64_t* qoff, uint64_t nq,
                          uint32_t k, uint32_t flags, uint32_t* start, uint32_t* end, uint64_t* hit_off)
    int sa_query_hits_r(sa_index* idx, sa_result* res, uint32_t flags, uint32_t* hit_pos, uint32_t* hit_doc, uint64_t cap,
                        uint64_t* hit_off_out)
    int sa_query_records_size(sa_index* idx, sa_result* res, uint32_t flags, uint64_t* num_records, uint64_t* num_bytes)
    int sa_query_records(sa_index* idx, sa_result* res, uint8_t* out, uint64_t cap_bytes, uint64_t* rec_off,
                         uint64_t cap_records, uint64_t* hit_off_out, uint32_t* hit_doc)
    int sa_attach_csv_file(sa_index* idx, const char* csv_path)
    int sa_build_text_multi(sa_index** idxs, int ndev, const uint8_t* text, uint64_t n, uint32_t flags)
    int sa_save(sa_index* idx, const char* path)
    int sa_load(sa_index* idx, const char* path)

cdef enum:
    SA_FLAG_LOWERCASE = 1
    SA_QUERY_FILE_SEMANTICS = 1
    SA_QUERY_DEDUPE = 2
    SA_RECORDS_REFERENCE_COMPAT = 4
    MAX_COLUMNS = 64


class SaCudaError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"libsa_cuda error {code}: {msg}")
        self.code = code


cdef int _check(int rc) except -1:
    if rc != 0:
        raise SaCudaError(rc, sa_last_error().decode("utf-8", "replace"))
    return 0


cdef class _Handle:
    """One ``sa_index*`` (a partition of the documents, or one search column of a CSV file)."""
    cdef sa_index* h
    cdef sa_result* r         # per-call query state (one batch at a time per handle)

    def __cinit__(self, uint32_t max_suffix_length, int device):
        self.h = NULL
        self.r = NULL
        _check(sa_create(&self.h, max_suffix_length, device))

    def __dealloc__(self):
        self.close()

    def close(self):
        if self.r != NULL:
            sa_result_destroy(self.r)
            self.r = NULL
        if self.h != NULL:
            sa_destroy(self.h)
            self.h = NULL

    def build_text(self, text, bint lowercase, uint64_t tail=0):
        """text: contiguous uint8 NumPy array (pyx:173-180: one partition of the joined documents; its
        last `tail` bytes are the look-ahead into the next partition)."""
        cdef uintptr_t p = text.ctypes.data if text.size else 0
        cdef uint64_t n = text.size - tail
        cdef uint32_t flags = SA_FLAG_LOWERCASE if lowercase else 0
        cdef int rc
        with nogil:
            rc = sa_build_text_tail(self.h, <const uint8_t*> p, n, tail, flags)
        _check(rc)

    def build_csv(self, str path, str column):
        cdef bytes b_path = path.encode("utf-8")
        cdef bytes b_col = column.encode("utf-8")
        cdef const char* c_path = b_path
        cdef const char* c_col = b_col
        cdef int rc
        with nogil:
            rc = sa_build_csv(self.h, c_path, c_col, 0)
        _check(rc)

    def build_csv_partition(self, str path, str column, uint64_t partition_start=0):
        """One 2 GiB partition of a big file (engine.c:1469-1481) -> (partition_end, file_bytes)."""
        cdef bytes b_path = path.encode("utf-8")
        cdef bytes b_col = column.encode("utf-8")
        cdef const char* c_path = b_path
        cdef const char* c_col = b_col
        cdef uint64_t end = 0, size = 0
        cdef int rc
        with nogil:
            rc = sa_build_csv_partition(self.h, c_path, c_col, partition_start, 0, &end, &size)
        _check(rc)
        return int(end), int(size)

    def attach_csv_file(self, str path):
        _check(sa_attach_csv_file(self.h, path.encode("utf-8")))

    @property
    def n(self):
        return int(sa_size(self.h))

    def sa(self):
        n = self.n
        out = np.empty(max(n, 1), dtype=np.uint32)
        cdef uintptr_t p = out.ctypes.data
        _check(sa_copy_sa_u32(self.h, <uint32_t*> p, n))
        return out[:n]

    def text(self):
        n = self.n
        out = np.empty(max(n, 1), dtype=np.uint8)
        cdef uintptr_t p = out.ctypes.data
        _check(sa_copy_text(self.h, <uint8_t*> p, n))
        return out[:n]

    def save(self, str path):
        _check(sa_save(self.h, path.encode()))

    def load(self, str path):
        _check(sa_load(self.h, path.encode()))

    cdef _ranges(self, list queries, uint32_t k, uint32_t flags):
        if self.r == NULL:
            _check(sa_result_create(self.h, &self.r))
        cdef uint64_t nq = len(queries)
        lens = np.fromiter((len(q) for q in queries), dtype=np.uint64, count=nq)
        qoff = np.zeros(nq + 1, dtype=np.uint64)
        np.cumsum(lens, out=qoff[1:])
        joined = b"".join(queries)
        qbytes = np.frombuffer(joined, dtype=np.uint8) if joined else np.zeros(1, dtype=np.uint8)
        start = np.empty(max(nq, 1), dtype=np.uint32)
        end = np.empty(max(nq, 1), dtype=np.uint32)
        hoff = np.zeros(nq + 1, dtype=np.uint64)
        cdef uintptr_t pq = qbytes.ctypes.data, po = qoff.ctypes.data
        cdef uintptr_t ps = start.ctypes.data, pe = end.ctypes.data, ph = hoff.ctypes.data
        cdef int rc
        with nogil:
            rc = sa_query_ranges_r(self.h, self.r, <const uint8_t*> pq, <const uint64_t*> po, nq, k, flags, <uint32_t*> ps,
                                   <uint32_t*> pe, <uint64_t*> ph)
        _check(rc)
        return start[:nq], end[:nq], hoff

    def query(self, list queries, uint32_t k, bint file_semantics=False, bint dedupe=False):
        """queries: list[bytes] -> (start u32, end u32, hit_off u64[nq+1], hit_pos u32, hit_doc u32)."""
        start, end, hoff = self._ranges(queries, k, SA_QUERY_FILE_SEMANTICS if file_semantics else 0)
        cdef uint64_t nq = len(queries)
        cdef uint64_t total = int(hoff[nq])
        pos = np.empty(max(total, 1), dtype=np.uint32)
        doc = np.empty(max(total, 1), dtype=np.uint32)
        off = np.zeros(nq + 1, dtype=np.uint64)
        cdef uintptr_t pp = pos.ctypes.data, pd = doc.ctypes.data, pf = off.ctypes.data
        cdef uint32_t flags = SA_QUERY_DEDUPE if dedupe else 0
        cdef int rc
        with nogil:
            rc = sa_query_hits_r(self.h, self.r, flags, <uint32_t*> pp, <uint32_t*> pd, total, <uint64_t*> pf)
        _check(rc)
        out = int(off[nq]) if nq else 0
        return start, end, off, pos[:out], doc[:out]

    def query_records(self, list queries, uint32_t k, bint file_semantics=False, bint dedupe=False, bint compat=False):
        """-> (hit_off u64[nq+1], rec_off u64[records+1], bytes u8, record ids u32): the records of the
        first min(k, count) hits of every query, cut and gathered on the device."""
        self._ranges(queries, k, SA_QUERY_FILE_SEMANTICS if file_semantics else 0)
        cdef uint64_t nq = len(queries)
        cdef uint64_t num = 0, nbytes = 0
        cdef uint32_t flags = (SA_QUERY_DEDUPE if dedupe else 0) | (SA_RECORDS_REFERENCE_COMPAT if compat else 0)
        cdef int rc
        with nogil:
            rc = sa_query_records_size(self.h, self.r, flags, &num, &nbytes)
        _check(rc)
        buf = np.empty(max(nbytes, 1), dtype=np.uint8)
        roff = np.zeros(num + 1, dtype=np.uint64)
        hoff = np.zeros(nq + 1, dtype=np.uint64)
        ids = np.empty(max(num, 1), dtype=np.uint32)
        cdef uintptr_t pb = buf.ctypes.data, pr = roff.ctypes.data, ph = hoff.ctypes.data, pi = ids.ctypes.data
        with nogil:
            rc = sa_query_records(self.h, self.r, <uint8_t*> pb, nbytes, <uint64_t*> pr, num, <uint64_t*> ph, <uint32_t*> pi)
        _check(rc)
        return hoff, roff, buf[:nbytes], ids[:num]


def _build_csv_columns(list handles, str path_s, list columns):
    """handles[c] <- column columns[c] of the file: one read / upload / quote analysis for all."""
    cdef sa_index* hs[MAX_COLUMNS]
    cdef const char* cs[MAX_COLUMNS]
    cdef uint32_t n = len(handles)
    if n == 0 or n > MAX_COLUMNS or n != len(columns):
        raise ValueError("one handle per column, at most 64 columns")
    enc = [c.encode("utf-8") for c in columns]          # keeps the byte strings alive during the call
    cdef uint32_t i
    cdef bytes name
    for i in range(n):
        hs[i] = (<_Handle> handles[i]).h
        name = enc[i]
        cs[i] = name                                     # `enc` keeps every name alive
    cdef bytes path = path_s.encode("utf-8")
    cdef const char* c_path = path
    cdef int rc
    with nogil:
        rc = sa_build_csv_columns(hs, cs, n, c_path, 0)
    _check(rc)


def _build_text_multi(list handles, text, bint lowercase):
    """ONE index over the GPUs of `handles` (sample sort, a thread per handle inside the library)."""
    cdef sa_index* hs[16]
    cdef int n = len(handles)
    if n < 1 or n > 16:
        raise ValueError("1..16 handles")
    cdef int i
    for i in range(n):
        hs[i] = (<_Handle> handles[i]).h
    text = np.ascontiguousarray(text, dtype=np.uint8)
    cdef uintptr_t p = text.ctypes.data if text.size else 0
    cdef uint64_t nbytes = text.size
    cdef uint32_t flags = SA_FLAG_LOWERCASE if lowercase else 0
    cdef int rc
    with nogil:
        rc = sa_build_text_multi(hs, n, <const uint8_t*> p, nbytes, flags)
    _check(rc)


class SuffixArray(SuffixArrayHost):
    """The reference's class surface (suffix_array.pyx:110-267); the logic is _host.SuffixArrayHost,
    shared with the ctypes class, bound here to the compiled handle type."""
    _Index = _Handle
    _build_csv_columns = staticmethod(_build_csv_columns)
    _build_text_multi = staticmethod(_build_text_multi)
