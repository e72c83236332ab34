"""Host logic of the ``SuffixArray`` class -- ONE copy, shared by the Cython class
(``suffix_array_cy.pyx``, the reference's host language) and the ctypes class (``suffix_array.py``).
The two differ only in how they reach ``libsa_cuda.so``: each supplies an index type with the small
method set listed under ``Backend`` below.

Mirrors ``suffix_array/suffix_array.pyx:110-267`` of the reference: same names, argument meaning and
error behaviour.  This layer joins / encodes strings, walks partitions and turns the packed record
buffer the library returns into Python objects; all index arithmetic -- build, bisection, hit
expansion, record ids, dedupe, record cutting -- runs in the CUDA library.

Backend (duck-typed): ``Index(max_suffix_length, device)`` with
``build_text(u8 array, lowercase, tail)``, ``build_csv(path, column)``, ``sa()``, ``text()``, ``n``,
``save(path)``, ``load(path)``, ``attach_csv_file(path)``, ``close()``,
``query(queries, k, file_semantics, dedupe) -> (start, end, hit_off, hit_pos, hit_doc)``,
``query_records(queries, k, file_semantics, dedupe, compat) -> (hit_off, rec_off, bytes, rec_ids)``;
and a module-level ``build_csv_columns(indexes, path, columns)``.
"""
from __future__ import annotations

import csv
import json
import os

import numpy as np


class SuffixArrayHost:
    # The reference cuts the joined text every 2 GiB into independent partitions, each with its own
    # suffix array, and a query walks them in order until k records are found
    # (suffix_array.pyx:148-180, :221-247).  Same here; tests shrink this to exercise the walk.
    PARTITION_BYTES = 1 << 31

    # supplied by the concrete class
    _Index = None
    _build_csv_columns = None
    _build_text_multi = None     # (indexes, u8 text, lowercase) -> one index sharded over the indexes' GPUs

    def __init__(self, max_suffix_length: int = 64, documents=None, csv_file: str | None = None,
                 search_column=None, device: int = 0, csv_query: str = "file", dedupe: bool = False,
                 reference_records: bool = False, devices=None):
        """csv_query: "file" = the reference's CSV bisection (get_substring_positions_file,
        engine.c:920-999: probes compare the raw file bytes -- the only query its Python class
        reaches, suffix_array.pyx:221-247) or "text" = get_substring_positions on the parsed
        column (engine.c:869-918).  dedupe: every record at most once per query (first hit wins; the
        reference keeps duplicates, engine.c:1187).  reference_records: cut CSV records the way
        get_matching_records_file does (last byte dropped, engine.c:1314; first byte too for hits at
        file offsets >= 4096).  devices=[0, 1, ...]: ONE index built over several GPUs of this process
        (sample sort, sa_build_text_multi; no torch.distributed needed): every GPU ends up with a replica
        and query batches are split over them; texts up to 2^32 - 1025 bytes stay ONE index (no 2 GiB
        partitioning).  A CSV index is built once per device (replicas)."""
        if csv_query not in ("file", "text"):
            raise ValueError('csv_query must be "file" or "text"')
        self.max_suffix_length = int(max_suffix_length)
        self._devices = [int(d) for d in devices] if devices else None
        if self._devices is not None and len(set(self._devices)) != len(self._devices):
            raise ValueError("devices must be distinct")
        self._device = self._devices[0] if self._devices else int(device)
        self._replicas = []          # devices=: the other GPUs' copies of the (single) index
        self.csv_query = csv_query
        self.dedupe = bool(dedupe)
        self.reference_records = bool(reference_records)
        self._parts = [self._Index(self.max_suffix_length, self._device)]   # one index per partition / search column
        self._mode = None            # "documents" | "csv"
        self.columns = None          # csv: header cells (pyx:189-194)
        self.search_columns = None   # csv: the indexed columns, in query order
        self._csv_path = None
        if documents is not None and csv_file is not None:
            raise ValueError("pass either documents= or csv_file=, not both")
        if documents is not None:
            self.construct_truncated_suffix_array_documents(documents)
        elif csv_file is not None:
            if search_column is None:
                raise ValueError("csv_file= needs search_column=")
            self.construct_truncated_suffix_array_from_csv(csv_file, search_column)

    def _drop_replicas(self) -> None:
        for r in self._replicas:
            r.close()
        self._replicas = []

    def _resize(self, count: int) -> None:
        for old in self._parts[count:]:
            old.close()
        self._parts = self._parts[:count] + [self._Index(self.max_suffix_length, self._device)
                                             for _ in range(count - len(self._parts))]

    # ------------------------------------------------------------------ builds
    def construct_truncated_suffix_array_documents(self, documents) -> None:
        """pyx:129-180: '\\n'.join(documents) -> utf-8 -> ASCII lower-case -> one index per 2 GiB."""
        if not isinstance(documents, list):
            try:
                documents = list(documents)                       # pyx:130-134
            except Exception:
                raise ValueError("Documents must be a list of strings")
        joined = "\n".join(documents).encode("utf-8")             # pyx:143 (no trailing '\n')
        text = np.frombuffer(joined, dtype=np.uint8)
        self._drop_replicas()
        if self._devices and len(self._devices) > 1:
            # one index over all the GPUs: every handle ends up with the text and the complete array
            self._resize(1)
            self._replicas = [self._Index(self.max_suffix_length, d) for d in self._devices[1:]]
            type(self)._build_text_multi([self._parts[0]] + self._replicas, text, True)
            self._text = self._doc_starts = None
            self._mode = "documents"
            return
        pb = int(self.PARTITION_BYTES)
        nparts = max(1, -(-text.size // pb))                      # pyx:148-152
        self._resize(nparts)
        look = 960                                                # sa_build_text_tail keeps at most this much
        for i, part in enumerate(self._parts):                    # pyx:173-180: cut at fixed byte offsets
            # the reference sorts partition i on `text + i * 2 GiB` of the ONE joined buffer: the bytes after
            # the cut are visible to the last suffixes of the partition -> pass them as look-ahead tail;
            # lowercase_string (pyx:103-107) runs on the device copy: SA_FLAG_LOWERCASE
            end = min(text.size, (i + 1) * pb)
            tail = min(look, text.size - end)
            part.build_text(np.ascontiguousarray(text[i * pb: end + tail]), True, tail)
        self._text = self._doc_starts = None
        if nparts > 1:
            # a document may straddle a cut: records of a partitioned text are cut from a host copy of the
            # joined text by GLOBAL position (a single partition -- any text below 2 GiB -- needs no host copy:
            # the library cuts the records on the device)
            low = text.copy()
            up = (low >= 65) & (low <= 90)
            low[up] += 32
            self._text = low.tobytes()
            self._doc_starts = np.concatenate([[0], np.flatnonzero(low == 10).astype(np.int64) + 1])
        self._mode = "documents"

    def construct_truncated_suffix_array_from_csv(self, filename: str, search_column) -> None:
        """pyx:183-207 -> engine.c:1454: scan of the search column(s), build, remap to file offsets.
        search_column: one name, or a list of names (the reference's C entry takes `char** search_cols`
        but indexes the first only, engine.c:1476; pyx:197 "TODO: Allow multiple columns"): one index
        per column from ONE read of the file."""
        if not os.path.exists(filename):
            raise FileNotFoundError(filename)
        with open(filename, "r", newline="") as f:                # pyx:189-194
            self.columns = f.readline().strip().split(",")
        cols = [search_column] if isinstance(search_column, str) else list(search_column)
        if not cols or len(set(cols)) != len(cols):
            raise ValueError("search_column must name one or more distinct columns")
        self._drop_replicas()
        if os.path.getsize(filename) >= self.PARTITION_BYTES:
            # engine.c:1437-1438, :1469-1481: the FILE is cut every 2 GiB, one suffix array per partition, each
            # partition resuming where the previous one stopped; queries walk them in order (pyx:221-247)
            if len(cols) != 1:
                raise ValueError("files of 2 GiB and more are indexed one search column at a time")
            self._resize(1)
            start, size, p = 0, None, 0
            while size is None or start < size:
                if p >= len(self._parts):
                    self._parts.append(self._Index(self.max_suffix_length, self._device))
                start, size = self._parts[p].build_csv_partition(filename, cols[0], start)
                p += 1
            self._resize(p)
            self.search_columns = cols
            self._csv_path = filename
            self._mode = "csv"
            self._csv_partitions = p
            return
        self._csv_partitions = 1
        self._resize(len(cols))
        if len(cols) == 1:
            self._parts[0].build_csv(filename, cols[0])
            if self._devices and len(self._devices) > 1:          # replicas: the same build on every other GPU
                import threading
                self._replicas = [self._Index(self.max_suffix_length, d) for d in self._devices[1:]]
                errs = []

                def _one(ix):
                    try:
                        ix.build_csv(filename, cols[0])
                    except Exception as ex:                       # noqa: BLE001
                        errs.append(ex)
                th = [threading.Thread(target=_one, args=(ix,)) for ix in self._replicas]
                for t in th:
                    t.start()
                for t in th:
                    t.join()
                if errs:
                    raise errs[0]
        else:
            type(self)._build_csv_columns(self._parts, filename, cols)
        self.search_columns = cols
        self._csv_path = filename
        self._mode = "csv"

    # ------------------------------------------------------------------ queries
    def query_records(self, substring: str, k: int = 1000, column: str | None = None):
        if substring == "":                                        # pyx:210-211
            return []
        return self.query_records_batch([substring], k, column=column)[0]

    def _walk(self, column):
        if self._mode == "csv" and column is not None and getattr(self, "_csv_partitions", 1) == 1:
            if column not in self.search_columns:
                raise ValueError(f"column {column!r} is not indexed (indexed: {self.search_columns})")
            return [self._parts[self.search_columns.index(column)]]
        return list(self._parts)

    def query_records_batch(self, substrings, k: int = 1000, column: str | None = None):
        """One list of records per query: the first min(k, matches) hits in suffix-array order
        (engine.c:1185-1211 / :1356-1388), duplicates kept unless dedupe=True; partitions (documents)
        or search columns (csv) are walked in order until k records are found (pyx:221-247).
        Documents index: the matching documents (lower-cased, as indexed); CSV index: one
        dict(zip(header, row)) per hit (pyx:258-260).  The records are cut on the device and arrive
        as one packed buffer (sa_query_records)."""
        if self._mode is None:
            raise RuntimeError("no suffix array has been constructed")
        k = int(k)
        qs = [s.lower().encode("utf-8") for s in substrings]      # pyx:228
        nonempty = [i for i, q in enumerate(qs) if q]
        out = [[] for _ in qs]
        if not nonempty:
            return out
        sub = [qs[i] for i in nonempty]
        is_csv = self._mode == "csv"
        file_sem = is_csv and self.csv_query == "file"
        if not is_csv and len(self._parts) > 1:
            return self._query_partitioned(sub, nonempty, out, k)
        if self._replicas and len(self._parts) == 1 and len(sub) >= 2 * (1 + len(self._replicas)):
            return self._query_replicas(sub, nonempty, out, k, is_csv, file_sem)
        recs = [[] for _ in sub]
        seen = [set() for _ in sub] if (self.dedupe and len(self._walk(column)) > 1) else None
        for p, part in enumerate(self._walk(column)):
            need = [j for j in range(len(sub)) if len(recs[j]) < k]
            if not need:
                break
            kk = max(k - len(recs[j]) for j in need)
            hoff, roff, buf, ids = part.query_records([sub[j] for j in need], kk, file_sem, self.dedupe,
                                                      is_csv and self.reference_records)
            data = buf.tobytes()
            ro = roff.tolist()
            for t, j in enumerate(need):
                a, b = int(hoff[t]), int(hoff[t + 1])
                for h in range(a, b):
                    if len(recs[j]) >= k:
                        break
                    if seen is not None:                          # dedupe across partitions / columns
                        key = (p if not is_csv else 0, int(ids[h]))
                        if key in seen[j]:
                            continue
                        seen[j].add(key)
                    recs[j].append(data[ro[h]:ro[h + 1]])
        for j, i in enumerate(nonempty):
            out[i] = self._decode(recs[j], is_csv)
        return out

    def _decode(self, recs, is_csv):
        if is_csv:
            lines = [r.decode("utf-8", "replace").rstrip("\r") for r in recs]
            return [dict(zip(self.columns, row)) for row in csv.reader(lines)]   # pyx:258-260
        return [r.decode("utf-8", "replace") for r in recs]

    def _query_replicas(self, sub, nonempty, out, k, is_csv, file_sem):
        """devices=[...]: every GPU holds a replica of the index; the batch is split over them (no
        collective: SURVEY 8(e) "replicas / batch split")."""
        import threading
        reps = [self._parts[0]] + self._replicas
        cut = [len(sub) * g // len(reps) for g in range(len(reps) + 1)]
        res, errs = [None] * len(reps), []

        def _one(g):
            try:
                res[g] = reps[g].query_records(sub[cut[g]:cut[g + 1]], k, file_sem, self.dedupe,
                                               is_csv and self.reference_records)
            except Exception as ex:                               # noqa: BLE001
                errs.append(ex)
        th = [threading.Thread(target=_one, args=(g,)) for g in range(len(reps))]
        for t in th:
            t.start()
        for t in th:
            t.join()
        if errs:
            raise errs[0]
        for g in range(len(reps)):
            hoff, roff, buf, _ = res[g]
            data, ro = buf.tobytes(), roff.tolist()
            for t in range(cut[g + 1] - cut[g]):
                recs = [data[ro[h]:ro[h + 1]] for h in range(int(hoff[t]), int(hoff[t + 1]))]
                out[nonempty[cut[g] + t]] = self._decode(recs, is_csv)
        return out

    def _query_partitioned(self, sub, nonempty, out, k):
        """documents cut into several partitions (pyx:148-180): walk them in order until k records are
        found (pyx:221-247); a hit's record is the document around its GLOBAL text position."""
        pb = int(self.PARTITION_BYTES)
        ds, text = self._doc_starts, self._text
        docs = [[] for _ in sub]
        for p, part in enumerate(self._parts):
            need = [j for j in range(len(sub)) if len(docs[j]) < k]
            if not need:
                break
            kk = max(k - len(docs[j]) for j in need)
            _, _, off, pos, _ = part.query([sub[j] for j in need], kk, False, False)
            for t, j in enumerate(need):
                gpos = pos[int(off[t]):int(off[t + 1])].astype(np.int64) + p * pb
                for d in (np.searchsorted(ds, gpos, side="right") - 1).tolist():
                    if len(docs[j]) >= k:
                        break
                    if self.dedupe and d in docs[j]:
                        continue
                    docs[j].append(d)
        for j, i in enumerate(nonempty):
            recs = []
            for d in docs[j]:
                a = int(ds[d])
                b = int(ds[d + 1]) - 1 if d + 1 < ds.size else len(text)
                recs.append(text[a:b].decode("utf-8", "replace"))
            out[i] = recs
        return out

    def query_positions(self, substrings, k: int = 1000):
        """-> (start, end, hit_off, hit_pos, hit_doc) NumPy arrays straight from the C ABI (first
        partition / column): inclusive rank range with the reference's UINT32_MAX sentinels, positions
        (file offsets on the CSV path) and record ids of the first min(k, count) hits."""
        qs = [s.lower().encode("utf-8") if isinstance(s, str) else bytes(s) for s in substrings]
        return self._parts[0].query(qs, int(k), self._mode == "csv" and self.csv_query == "file", self.dedupe)

    # ------------------------------------------------------------------ persistence
    def save(self, path: str) -> None:
        """Write the built index to `path` (+ `path.partNNN` for further partitions / columns, +
        `path.json` with what the host side needs).  The reference only sketches this
        (engine.c:1098-1165, TODO.md)."""
        if self._mode is None:
            raise RuntimeError("no suffix array has been constructed")
        for i, idx in enumerate(self._parts):
            idx.save(path if i == 0 else f"{path}.part{i:03d}")
        meta = {"mode": self._mode, "max_suffix_length": self.max_suffix_length, "partitions": len(self._parts),
                "partition_bytes": int(self.PARTITION_BYTES), "columns": self.columns, "csv_path": self._csv_path,
                "search_columns": self.search_columns, "csv_query": self.csv_query, "dedupe": self.dedupe,
                "reference_records": self.reference_records}
        with open(path + ".json", "w") as f:
            json.dump(meta, f)

    @classmethod
    def load(cls, path: str, device: int = 0):
        """Restore an index written by save(): no sorting, the arrays go straight to the device; a CSV
        index re-attaches its file (records and file-semantics probes read it)."""
        with open(path + ".json") as f:
            meta = json.load(f)
        self = cls(max_suffix_length=int(meta["max_suffix_length"]), device=device,
                   csv_query=meta.get("csv_query", "file"), dedupe=bool(meta.get("dedupe", False)),
                   reference_records=bool(meta.get("reference_records", False)))
        self.PARTITION_BYTES = int(meta["partition_bytes"])
        self._resize(int(meta["partitions"]))
        for i, idx in enumerate(self._parts):
            idx.load(path if i == 0 else f"{path}.part{i:03d}")
        self._mode = meta["mode"]
        self._text = self._doc_starts = None
        if self._mode == "documents" and len(self._parts) > 1:
            low = np.concatenate([idx.text() for idx in self._parts])
            self._text = low.tobytes()
            self._doc_starts = np.concatenate([[0], np.flatnonzero(low == 10).astype(np.int64) + 1])
        if self._mode == "csv":
            self.columns = meta["columns"]
            self.search_columns = meta["search_columns"]
            self._csv_path = meta["csv_path"]
            for idx in self._parts:
                idx.attach_csv_file(self._csv_path)
        return self

    # ------------------------------------------------------------------ misc
    @property
    def suffix_array(self):
        """uint32 array as SuffixArray_struct.suffix_array holds it (file offsets on the CSV path);
        a list of arrays when the index has several partitions / search columns."""
        arrays = [idx.sa() for idx in self._parts]
        return arrays[0] if len(arrays) == 1 else arrays

    def close(self) -> None:
        for idx in self._parts:
            idx.close()
        self._parts = []
        self._drop_replicas()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
