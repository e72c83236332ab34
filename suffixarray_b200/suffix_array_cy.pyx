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
``suffixarray_b200/suffix_array.py`` is the same class over ctypes (no compiler needed); the two are
run through the same drop-in tests.  Built by ``__graft_entry__.build()`` (cython + gcc, linked
against ``libsa_cuda.so`` with an ``$ORIGIN`` rpath).
"""
from libc.stdint cimport uint8_t, uint32_t, uint64_t, uintptr_t

import csv
import json
import mmap
import os

import numpy as np

cdef extern from "sa_cuda.h" nogil:
    ctypedef struct sa_index:
        pass
    const char* sa_last_error()
    int sa_create(sa_index** out, uint32_t max_suffix_length, int device)
    void sa_destroy(sa_index* idx)
    int sa_build_text(sa_index* idx, const uint8_t* text, uint64_t n, uint32_t flags)
    int sa_build_csv(sa_index* idx, const char* csv_path, const char* column, uint32_t flags)
    int sa_build_csv_columns(sa_index** idxs, const char** columns, uint32_t ncols, const char* csv_path, uint32_t flags)
    uint64_t sa_size(const sa_index* idx)
    int sa_copy_sa_u32(sa_index* idx, uint32_t* out, uint64_t cap)
    int sa_copy_text(sa_index* idx, uint8_t* out, uint64_t cap)
    int sa_query_ranges(sa_index* idx, const uint8_t* qbytes, const uint64_t* qoff, uint64_t nq, uint32_t k,
                        uint32_t* start, uint32_t* end, uint64_t* hit_off)
    int sa_query_hits(sa_index* idx, uint32_t* hit_pos, uint32_t* hit_doc, uint64_t cap)
    int sa_save(sa_index* idx, const char* path)
    int sa_load(sa_index* idx, const char* path)

cdef enum:
    SA_FLAG_LOWERCASE = 1
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

    def __cinit__(self, uint32_t max_suffix_length, int device):
        self.h = NULL
        _check(sa_create(&self.h, max_suffix_length, device))

    def __dealloc__(self):
        if self.h != NULL:
            sa_destroy(self.h)
            self.h = NULL

    def close(self):
        if self.h != NULL:
            sa_destroy(self.h)
            self.h = NULL

    def build_text(self, text, bint lowercase):
        """text: contiguous uint8 NumPy array (pyx:173-180: one partition of the joined documents)."""
        cdef uintptr_t p = text.ctypes.data if text.size else 0
        cdef uint64_t n = text.size
        cdef uint32_t flags = SA_FLAG_LOWERCASE if lowercase else 0
        cdef int rc
        with nogil:
            rc = sa_build_text(self.h, <const uint8_t*> p, n, flags)
        _check(rc)

    def build_csv(self, bytes path, bytes column):
        cdef const char* c_path = path
        cdef const char* c_col = column
        cdef int rc
        with nogil:
            rc = sa_build_csv(self.h, c_path, c_col, 0)
        _check(rc)

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

    def query(self, list queries, uint32_t k):
        """queries: list[bytes] -> (start u32, end u32, hit_off u64[nq+1], hit_pos u32, hit_doc u32)."""
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
            rc = sa_query_ranges(self.h, <const uint8_t*> pq, <const uint64_t*> po, nq, k, <uint32_t*> ps, <uint32_t*> pe,
                                 <uint64_t*> ph)
        _check(rc)
        cdef uint64_t total = int(hoff[nq])
        pos = np.empty(max(total, 1), dtype=np.uint32)
        doc = np.empty(max(total, 1), dtype=np.uint32)
        cdef uintptr_t pp = pos.ctypes.data, pd = doc.ctypes.data
        with nogil:
            rc = sa_query_hits(self.h, <uint32_t*> pp, <uint32_t*> pd, total)
        _check(rc)
        return start[:nq], end[:nq], hoff, pos[:total], doc[:total]


def _build_csv_columns(list handles, bytes path, list columns):
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
    cdef const char* c_path = path
    cdef int rc
    with nogil:
        rc = sa_build_csv_columns(hs, cs, n, c_path, 0)
    _check(rc)


cdef class SuffixArray:
    cdef public uint32_t max_suffix_length
    cdef public object columns              # csv: header cells (pyx:189-194)
    cdef public object search_columns       # csv: the indexed columns, in query order
    cdef public object PARTITION_BYTES      # pyx:148-152 cuts the joined text every 2 GiB
    cdef int _device
    cdef list _parts                        # documents: one handle per partition; csv: one per search column
    cdef object _mode                       # None | "documents" | "csv"
    cdef object _text, _doc_starts
    cdef object _csv_path, _csv_mm, _row_starts, _row_ends

    def __init__(self, max_suffix_length=64, documents=None, csv_file=None, search_column=None, device=0):
        self.max_suffix_length = int(max_suffix_length)
        self._device = int(device)
        self.PARTITION_BYTES = 1 << 31
        self._parts = [_Handle(self.max_suffix_length, self._device)]
        self._mode = None
        self.columns = None
        self.search_columns = None
        self._csv_mm = None
        if documents is not None and csv_file is not None:
            raise ValueError("pass either documents= or csv_file=, not both")
        if documents is not None:
            self.construct_truncated_suffix_array_documents(documents)
        elif csv_file is not None:
            if search_column is None:
                raise ValueError("csv_file= needs search_column=")
            self.construct_truncated_suffix_array_from_csv(csv_file, search_column)

    def _resize(self, int count):
        for old in self._parts[count:]:
            old.close()
        self._parts = self._parts[:count] + [_Handle(self.max_suffix_length, self._device)
                                             for _ in range(count - len(self._parts))]

    # ------------------------------------------------------------------ builds
    cpdef construct_truncated_suffix_array_documents(self, documents):
        """pyx:129-180: '\\n'.join(documents) -> utf-8 -> ASCII lower-case -> one index per 2 GiB."""
        if not isinstance(documents, list):
            try:
                documents = list(documents)                       # pyx:130-134
            except Exception:
                raise ValueError("Documents must be a list of strings")
        joined = "\n".join(documents).encode("utf-8")             # pyx:143 (no trailing '\n')
        text = np.frombuffer(joined, dtype=np.uint8)
        pb = int(self.PARTITION_BYTES)
        nparts = max(1, -(-text.size // pb))                      # pyx:148-152
        self._resize(nparts)
        for i, part in enumerate(self._parts):                    # pyx:173-180: cut at fixed byte offsets
            # lowercase_string (pyx:103-107) runs on the device copy: SA_FLAG_LOWERCASE
            (<_Handle> part).build_text(np.ascontiguousarray(text[i * pb:(i + 1) * pb]), True)
        low = text.copy()                                         # host copy of what was indexed
        up = (low >= 65) & (low <= 90)
        low[up] += 32
        self._text = low.tobytes()
        nl = np.flatnonzero(low == 10).astype(np.int64)
        self._doc_starts = np.concatenate([[0], nl + 1]).astype(np.int64)
        self._mode = "documents"

    cpdef construct_truncated_suffix_array_from_csv(self, filename, search_column):
        """pyx:183-207 -> engine.c:1454: scan of the search column(s), build, remap to file offsets."""
        if not os.path.exists(filename):
            raise FileNotFoundError(filename)
        with open(filename, "r", newline="") as f:                # pyx:189-194
            self.columns = f.readline().strip().split(",")
        cols = [search_column] if isinstance(search_column, str) else list(search_column)
        if not cols or len(set(cols)) != len(cols):
            raise ValueError("search_column must name one or more distinct columns")
        self._resize(len(cols))
        if len(cols) == 1:
            (<_Handle> self._parts[0]).build_csv(filename.encode("utf-8"), cols[0].encode("utf-8"))
        else:
            _build_csv_columns(self._parts, filename.encode("utf-8"), cols)
        self.search_columns = cols
        self._csv_path = filename
        self._open_csv(filename)
        self._mode = "csv"

    def _open_csv(self, filename):
        if self._csv_mm is not None:
            self._close_csv()
        f = open(filename, "rb")
        self._csv_mm = mmap.mmap(f.fileno(), 0, access=mmap.ACCESS_READ)
        f.close()
        data = np.frombuffer(self._csv_mm, dtype=np.uint8)
        # record boundaries: '\n' outside quotes (RFC 4180; the reference seeks for them per hit,
        # engine.c:1217-1323).  Row 0 is the header line, as in the reference's scan (engine.c:513).
        in_quotes = (np.cumsum(data == 34, dtype=np.int64) & 1).astype(bool)
        nl = np.flatnonzero((data == 10) & ~in_quotes)
        starts = np.concatenate([[0], nl + 1]).astype(np.int64)
        if starts[-1] >= data.size:
            starts = starts[:-1]
        self._row_starts = starts
        self._row_ends = np.concatenate([nl, [data.size]]).astype(np.int64)[: starts.size]
        del data, in_quotes

    def _close_csv(self):
        self._row_starts = self._row_ends = None
        if self._csv_mm is not None:
            try:
                self._csv_mm.close()
            except BufferError:
                pass
            self._csv_mm = None

    # ------------------------------------------------------------------ queries
    cpdef query_records(self, substring, k=1000, column=None):
        if substring == "":                                        # pyx:210-211
            return []
        return self.query_records_batch([substring], k, column)[0]

    def query_records_batch(self, substrings, k=1000, column=None):
        """One list of records per query: the first min(k, matches) hits in suffix-array order,
        duplicates kept (engine.c:1185-1211 / :1356-1388); partitions (documents) or search columns
        (csv) are walked in order until k records are found (pyx:221-247)."""
        if self._mode is None:
            raise RuntimeError("no suffix array has been constructed")
        k = int(k)
        qs = [s.lower().encode("utf-8") for s in substrings]      # pyx:228
        nonempty = [i for i, q in enumerate(qs) if q]
        out = [[] for _ in qs]
        if not nonempty:
            return out
        sub = [qs[i] for i in nonempty]
        if self._mode == "csv":
            if column is not None and column not in self.search_columns:
                raise ValueError(f"column {column!r} is not indexed (indexed: {self.search_columns})")
            walk = [self._parts[self.search_columns.index(column)]] if column is not None else list(self._parts)
        else:
            walk = list(self._parts)
        pb = int(self.PARTITION_BYTES)
        found = [np.zeros(0, np.int64) for _ in sub]
        for p, part in enumerate(walk):
            need = [j for j in range(len(sub)) if found[j].size < k]
            if not need:
                break
            kk = max([k - found[j].size for j in need])
            _, _, off, pos, doc = (<_Handle> part).query([sub[j] for j in need], kk)
            for t, j in enumerate(need):
                a, b = int(off[t]), int(off[t + 1])
                if self._mode == "csv":
                    ids = doc[a:b].astype(np.int64)               # row of the FILE per hit (header = row 0)
                elif len(walk) == 1:
                    ids = doc[a:b].astype(np.int64)
                else:                                             # record of the global text position
                    ids = np.searchsorted(self._doc_starts, pos[a:b].astype(np.int64) + p * pb, side="right") - 1
                found[j] = np.concatenate([found[j], ids[: k - found[j].size]])
        for j, i in enumerate(nonempty):
            out[i] = self._materialise(found[j])
        return out

    def _materialise(self, ids):
        if self._mode == "documents":
            ds, text = self._doc_starts, self._text
            recs = []
            for d in ids.tolist():
                a = int(ds[d])
                b = int(ds[d + 1]) - 1 if d + 1 < ds.size else len(text)
                recs.append(text[a:b].decode("utf-8", "replace"))
            return recs
        rows = []
        for d in ids.tolist():
            line = self._csv_mm[int(self._row_starts[d]): int(self._row_ends[d])].decode("utf-8", "replace")
            rows.append(line.rstrip("\r"))
        return [dict(zip(self.columns, row)) for row in csv.reader(rows)]   # pyx:258-260

    # ------------------------------------------------------------------ persistence (SURVEY 8(f)-3)
    def save(self, str path):
        if self._mode is None:
            raise RuntimeError("no suffix array has been constructed")
        for i, part in enumerate(self._parts):
            (<_Handle> part).save(path if i == 0 else f"{path}.part{i:03d}")
        meta = {"mode": self._mode, "max_suffix_length": int(self.max_suffix_length), "partitions": len(self._parts),
                "partition_bytes": int(self.PARTITION_BYTES), "columns": self.columns, "csv_path": self._csv_path,
                "search_columns": self.search_columns}
        with open(path + ".json", "w") as f:
            json.dump(meta, f)

    @classmethod
    def load(cls, str path, device=0):
        with open(path + ".json") as f:
            meta = json.load(f)
        cdef SuffixArray self = cls(max_suffix_length=int(meta["max_suffix_length"]), device=device)
        self.PARTITION_BYTES = int(meta["partition_bytes"])
        self._resize(int(meta["partitions"]))
        for i, part in enumerate(self._parts):
            (<_Handle> part).load(path if i == 0 else f"{path}.part{i:03d}")
        self._mode = meta["mode"]
        if self._mode == "documents":
            low = np.concatenate([(<_Handle> part).text() for part in self._parts])
            self._text = low.tobytes()
            nl = np.flatnonzero(low == 10).astype(np.int64)
            self._doc_starts = np.concatenate([[0], nl + 1]).astype(np.int64)
        else:
            self.columns = meta["columns"]
            self.search_columns = meta["search_columns"]
            self._csv_path = meta["csv_path"]
            self._open_csv(self._csv_path)
        return self

    # ------------------------------------------------------------------ misc
    @property
    def suffix_array(self):
        """uint32 array as SuffixArray_struct.suffix_array holds it (file offsets on the CSV path);
        a list of arrays when the index has several partitions / search columns."""
        arrays = [(<_Handle> part).sa() for part in self._parts]
        return arrays[0] if len(arrays) == 1 else arrays

    def close(self):
        for part in self._parts:
            part.close()
        self._parts = []
        self._close_csv()

    # no __dealloc__ here: the handles free themselves (_Handle.__dealloc__ -> sa_destroy); the
    # reference never frees its arrays (SURVEY 8(b))
