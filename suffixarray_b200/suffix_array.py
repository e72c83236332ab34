"""Host-side mirror of the reference's Cython ``SuffixArray`` class, re-pointed at libsa_cuda.so.

Same names, argument meaning and error behaviour as ``suffix_array/suffix_array.pyx:110-267``
(the reference toolchain for this layer is Cython; the image has it, but the class holds no
arithmetic of its own, so the mirror is plain Python over the C ABI -- INTEGRATION.md shows the
``cdef extern from "sa_cuda.h"`` block a maintainer would drop into the .pyx instead):

* ``SuffixArray(max_suffix_length=64)``                                  pyx:125
* ``SuffixArray(documents=docs, max_suffix_length=32)``                  README.md:23
* ``SuffixArray(csv_file=f, search_column=c, max_suffix_length=32)``     README.md:40-44
* ``construct_truncated_suffix_array_documents(documents)``              pyx:129-180
* ``construct_truncated_suffix_array_from_csv(filename, search_column)`` pyx:183-207
  (``search_column`` may also be a list of column names -- the reference's C entry takes
  ``char** search_cols`` but indexes the first only, engine.c:1476, pyx:197 "TODO: Allow multiple
  columns": one index per column, built from ONE read of the file; queries walk the columns in
  the order given until k rows are found, or search one with ``column=``)
* ``query_records(substring, k=1000)``                                   pyx:209-267
* ``query_records_batch(substrings, k=1000)``  (new: the batched entry the GPU path wants)

All index arithmetic (build, bisection, hit expansion, doc ids) runs in the CUDA library; this
file only joins / encodes strings and turns (hit position, record id) pairs back into records.
There is no CPU fallback: without the built library or a GPU every call raises.
"""
from __future__ import annotations

import csv
import mmap
import os

import numpy as np

from . import _lib as S


class SuffixArray:
    # The reference cuts the joined text every 2 GiB into independent partitions, each with its own
    # suffix array, and a query walks them in order until k records are found
    # (suffix_array.pyx:148-180, :221-247).  Same here; tests shrink this to exercise the walk.
    PARTITION_BYTES = 1 << 31

    def __init__(self, max_suffix_length: int = 64, documents=None, csv_file: str | None = None,
                 search_column: str | list | None = None, device: int = 0):
        self.max_suffix_length = int(max_suffix_length)
        self._device = device
        self._idx = S.Index(self.max_suffix_length, device=device)
        self._parts = [self._idx]    # one Index per partition; _idx is partition 0
        self._mode = None            # "documents" | "csv"
        self._text = None            # documents: the joined, lower-cased utf-8 bytes (host copy)
        self._doc_starts = None
        self.columns = None          # csv: header cells (pyx:189-194)
        self.search_columns = None   # csv: the indexed columns, in query order
        self._col_idx = {}           # csv: column name -> its Index
        self._csv_path = None
        self._csv_mm = None
        self._row_starts = None
        if documents is not None and csv_file is not None:
            raise ValueError("pass either documents= or csv_file=, not both")
        if documents is not None:
            self.construct_truncated_suffix_array_documents(documents)
        elif csv_file is not None:
            if search_column is None:
                raise ValueError("csv_file= needs search_column=")
            self.construct_truncated_suffix_array_from_csv(csv_file, search_column)

    # ------------------------------------------------------------------ builds
    def construct_truncated_suffix_array_documents(self, documents) -> None:
        """pyx:129-180: '\\n'.join(documents) -> utf-8 -> ASCII lower-case -> build."""
        if not isinstance(documents, list):
            try:
                documents = list(documents)                     # pyx:130-134
            except Exception:
                raise ValueError("Documents must be a list of strings")
        joined = "\n".join(documents).encode("utf-8")           # pyx:143 (no trailing '\n')
        text = np.frombuffer(joined, dtype=np.uint8)
        pb = int(self.PARTITION_BYTES)
        nparts = max(1, -(-text.size // pb))                    # pyx:148-152
        for old in self._parts[1:]:
            old.close()
        self._parts = [self._idx] + [S.Index(self.max_suffix_length, device=self._device) for _ in range(nparts - 1)]
        for i, idx in enumerate(self._parts):                   # pyx:173-180: cut at fixed byte offsets
            # lowercase_string (pyx:103-107) runs on the device copy: SA_FLAG_LOWERCASE
            idx.build_text(text[i * pb:(i + 1) * pb], lowercase=True)
        low = text.copy()                                       # host copy of what was indexed
        up = (low >= 65) & (low <= 90)
        low[up] += 32
        self._text = low.tobytes()
        nl = np.flatnonzero(low == 10).astype(np.int64)
        self._doc_starts = np.concatenate([[0], nl + 1]).astype(np.int64)
        self._mode = "documents"

    def construct_truncated_suffix_array_from_csv(self, filename: str, search_column) -> None:
        """pyx:183-207 -> engine.c:1454: scan of the search column(s), build, remap to file offsets.
        search_column: one name, or a list of names (one index per column, one read of the file)."""
        if not os.path.exists(filename):
            raise FileNotFoundError(filename)
        with open(filename, "r", newline="") as f:              # pyx:189-194
            self.columns = f.readline().strip().split(",")
        cols = [search_column] if isinstance(search_column, str) else list(search_column)
        if not cols or len(set(cols)) != len(cols):
            raise ValueError("search_column must name one or more distinct columns")
        for old in self._parts[1:]:
            old.close()
        self._parts = [self._idx] + [S.Index(self.max_suffix_length, device=self._device) for _ in cols[1:]]
        if len(cols) == 1:
            self._idx.build_csv(filename, cols[0])
        else:
            S.Index.build_csv_columns(self._parts, filename, cols)
        self.search_columns = cols
        self._col_idx = dict(zip(cols, self._parts))
        self._csv_path = filename
        self._open_csv(filename)
        self._mode = "csv"

    def _open_csv(self, filename: str) -> None:
        f = open(filename, "rb")
        self._csv_mm = mmap.mmap(f.fileno(), 0, access=mmap.ACCESS_READ)
        f.close()
        data = np.frombuffer(self._csv_mm, dtype=np.uint8)
        # record boundaries: '\n' outside quotes (RFC 4180; the reference seeks for them per hit,
        # engine.c:1217-1323).  Row 0 is the header line, as in the reference's scan (engine.c:513).
        in_quotes = (np.cumsum(data == ord('"'), dtype=np.int64) & 1).astype(bool)
        nl = np.flatnonzero((data == ord("\n")) & ~in_quotes)
        starts = np.concatenate([[0], nl + 1]).astype(np.int64)
        if starts[-1] >= data.size:
            starts = starts[:-1]
        self._row_starts = starts
        self._row_ends = np.concatenate([nl, [data.size]]).astype(np.int64)[: starts.size]

    # ------------------------------------------------------------------ queries
    def query_records(self, substring: str, k: int = 1000, column: str | None = None):
        if substring == "":                                      # pyx:210-211
            return []
        return self.query_records_batch([substring], k, column=column)[0]

    def query_records_batch(self, substrings, k: int = 1000, column: str | None = None):
        """One list of records per query: the first min(k, matches) hits in suffix-array order,
        duplicates kept (engine.c:1185-1211 / :1356-1388).  CSV index over several columns: the
        columns are searched in the order they were given until k rows are found (`column=` limits
        the search to one of them)."""
        if self._mode is None:
            raise RuntimeError("no suffix array has been constructed")
        qs = [s.lower().encode("utf-8") for s in substrings]    # pyx:228
        nonempty = [i for i, q in enumerate(qs) if q]
        out = [[] for _ in qs]
        if not nonempty:
            return out
        sub = [qs[i] for i in nonempty]
        if self._mode == "csv":
            if column is not None and column not in self._col_idx:
                raise ValueError(f"column {column!r} is not indexed (indexed: {self.search_columns})")
            walk = [self._col_idx[column]] if column is not None else list(self._parts)
            rows = [np.zeros(0, np.int64) for _ in sub]
            for idx in walk:                                     # same walk as over partitions, pyx:221-247
                need = [j for j in range(len(sub)) if rows[j].size < k]
                if not need:
                    break
                kk = max(int(k) - rows[j].size for j in need)
                _, _, off = idx.query_ranges([sub[j] for j in need], k=kk)
                _, row = idx.query_hits(int(off[-1]))            # row of the FILE per hit (header = row 0)
                for t, j in enumerate(need):
                    take = row[int(off[t]):int(off[t + 1])][: int(k) - rows[j].size].astype(np.int64)
                    rows[j] = np.concatenate([rows[j], take])
            for j, i in enumerate(nonempty):
                out[i] = self._materialise(rows[j])
            return out
        start, end, off = self._idx.query_ranges(sub, k=int(k))
        pos, doc = self._idx.query_hits(int(off[-1]))
        if len(self._parts) == 1:
            for j, i in enumerate(nonempty):
                out[i] = self._materialise(doc[int(off[j]):int(off[j + 1])])
            return out
        # several partitions: walk them in order until k records are found (pyx:221-247)
        pb = int(self.PARTITION_BYTES)
        docs = [self._docs_of(pos[int(off[j]):int(off[j + 1])].astype(np.int64)) for j in range(len(sub))]
        for p, idx in enumerate(self._parts[1:], start=1):
            need = [j for j in range(len(sub)) if len(docs[j]) < k]
            if not need:
                break
            kk = max(int(k) - len(docs[j]) for j in need)
            s2, e2, off2 = idx.query_ranges([sub[j] for j in need], k=kk)
            pos2, _ = idx.query_hits(int(off2[-1]))
            for t, j in enumerate(need):
                take = pos2[int(off2[t]):int(off2[t + 1])][: int(k) - len(docs[j])].astype(np.int64) + p * pb
                docs[j] = np.concatenate([docs[j], self._docs_of(take)])
        for j, i in enumerate(nonempty):
            out[i] = self._materialise(docs[j])
        return out

    def _docs_of(self, global_pos: np.ndarray) -> np.ndarray:
        """record index of text positions (documents path): number of '\n' before the position."""
        return np.searchsorted(self._doc_starts, global_pos, side="right") - 1

    def query_positions(self, substrings, k: int = 1000):
        """-> (start, end, hit_off, hit_pos, hit_doc) NumPy arrays straight from the C ABI."""
        qs = [s.lower().encode("utf-8") if isinstance(s, str) else bytes(s) for s in substrings]
        start, end, off = self._idx.query_ranges(qs, k=int(k))
        pos, doc = self._idx.query_hits(int(off[-1]))
        return start, end, off, pos, doc

    def _materialise(self, docs: np.ndarray):
        if self._mode == "documents":
            ds, text = self._doc_starts, self._text
            recs = []
            for d in docs.tolist():
                a = int(ds[d])
                b = int(ds[d + 1]) - 1 if d + 1 < ds.size else len(text)
                recs.append(text[a:b].decode("utf-8", "replace"))
            return recs
        rows = []
        for d in docs.tolist():
            line = self._csv_mm[int(self._row_starts[d]): int(self._row_ends[d])].decode("utf-8", "replace")
            rows.append(line.rstrip("\r"))
        return [dict(zip(self.columns, row)) for row in csv.reader(rows)]   # pyx:258-260

    # ------------------------------------------------------------------ persistence
    def save(self, path: str) -> None:
        """Write the built index to `path` (+ `path.partNNN` for further partitions, + `path.json`
        with what the host side needs).  The reference only sketches this (engine.c:1098-1165)."""
        if self._mode is None:
            raise RuntimeError("no suffix array has been constructed")
        import json
        for i, idx in enumerate(self._parts):
            idx.save(path if i == 0 else f"{path}.part{i:03d}")
        meta = {"mode": self._mode, "max_suffix_length": self.max_suffix_length, "partitions": len(self._parts),
                "partition_bytes": int(self.PARTITION_BYTES), "columns": self.columns, "csv_path": self._csv_path,
                "search_columns": self.search_columns}
        with open(path + ".json", "w") as f:
            json.dump(meta, f)

    @classmethod
    def load(cls, path: str, device: int = 0) -> "SuffixArray":
        """Restore an index written by save(): no sorting, the arrays go straight to the device."""
        import json
        with open(path + ".json") as f:
            meta = json.load(f)
        self = cls(max_suffix_length=int(meta["max_suffix_length"]), device=device)
        self.PARTITION_BYTES = int(meta["partition_bytes"])
        self._parts = [self._idx] + [S.Index(self.max_suffix_length, device=device)
                                     for _ in range(int(meta["partitions"]) - 1)]
        for i, idx in enumerate(self._parts):
            idx.load(path if i == 0 else f"{path}.part{i:03d}")
        self._mode = meta["mode"]
        if self._mode == "documents":
            low = np.concatenate([idx.text() for idx in self._parts]) if self._parts else np.zeros(0, np.uint8)
            self._text = low.tobytes()
            nl = np.flatnonzero(low == 10).astype(np.int64)
            self._doc_starts = np.concatenate([[0], nl + 1]).astype(np.int64)
        else:
            self.columns = meta["columns"]
            self.search_columns = meta["search_columns"]
            self._col_idx = dict(zip(self.search_columns, self._parts))
            self._csv_path = meta["csv_path"]
            self._open_csv(self._csv_path)          # rows are materialised from the CSV file itself
        return self

    # ------------------------------------------------------------------ misc
    @property
    def suffix_array(self) -> np.ndarray:
        """uint32 array as SuffixArray_struct.suffix_array holds it (file offsets on the CSV path);
        a list of arrays when the index has several partitions / search columns."""
        return self._idx.sa() if len(self._parts) == 1 else [idx.sa() for idx in self._parts]

    def stats(self) -> dict:
        return self._idx.stats()

    def close(self) -> None:
        if self._idx is not None:
            for idx in self._parts:
                idx.close()
            self._parts = []
            self._idx = None
        if self._csv_mm is not None:
            self._row_starts = self._row_ends = None
            try:
                self._csv_mm.close()
            except BufferError:
                pass
            self._csv_mm = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
