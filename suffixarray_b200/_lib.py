"""ctypes binding of libsa_cuda.so (the C ABI declared in include/sa_cuda.h).

The library is built in-tree by ``suffixarray_b200/csrc/Makefile`` (``__graft_entry__.build()``).
There is no CPU fallback: if the shared object is missing, or no CUDA device is present,
every entry point raises.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsa_cuda.so")
CSRC = os.path.join(_HERE, "csrc")

SA_OK = 0
SA_FLAG_LOWERCASE = 1
SA_FLAG_TIMING = 2
SA_FLAG_DEVICE_TEXT = 4
SA_FLAG_SHARDED_SA = 8
SA_QUERY_FILE_SEMANTICS = 1
SA_QUERY_DEDUPE = 2
SA_RECORDS_REFERENCE_COMPAT = 4
UINT32_MAX = 0xFFFFFFFF

# every symbol include/sa_cuda.h declares (checked by tests/test_abi.py)
EXPORTS = [
    "sa_last_error", "sa_abi_version", "sa_create", "sa_destroy", "sa_build_text", "sa_build_text_tail", "sa_build_text_device",
    "sa_build_csv", "sa_build_csv_partition", "sa_build_csv_columns", "sa_size", "sa_copy_sa_u32", "sa_copy_sa_i64", "sa_copy_text_positions_u32",
    "sa_copy_quoted_bits", "sa_copy_text", "sa_device_sa", "sa_get_stats", "sa_query_ranges", "sa_query_hits",
    "sa_query_ranges_device", "sa_num_docs", "sa_copy_doc_starts", "sa_debug_copy_lcp", "sa_debug_set",
    "sa_shard_begin", "sa_shard_begin_csv", "sa_shard_text_size", "sa_shard_copy_text", "sa_shard_pack", "sa_shard_split", "sa_shard_sort", "sa_shard_edges", "sa_shard_finish",
    "sa_adopt_sa", "sa_adopt_sa_parts", "sa_shard_count", "sa_shard_recv_prepare", "sa_shard_scatter_p2p", "sa_shard_sort_recv",
    "sa_shard_final_prepare", "sa_shard_finish_bcast", "sa_shard_mark_built",
    "sa_save", "sa_load", "sa_export_reference_files",
    "sa_result_create", "sa_result_destroy", "sa_query_ranges_r", "sa_query_hits_r", "sa_query_records_size",
    "sa_query_records", "sa_result_stats", "sa_num_rows", "sa_copy_row_starts", "sa_attach_csv_file",
    "sa_comm_create_local", "sa_comm_create_shm", "sa_comm_destroy", "sa_comm_rank", "sa_comm_size", "sa_comm_barrier",
    "sa_shard_build", "sa_shard_info", "sa_build_text_multi", "sa_hit_offsets_noalloc",
]
SA_NOALLOC_MAX_QUERIES = 1 << 24


class SaStats(C.Structure):
    _fields_ = [
        ("n", C.c_uint64), ("max_suffix_length", C.c_uint32), ("lprime", C.c_uint32), ("rounds", C.c_uint32),
        ("radix_passes", C.c_uint32), ("kernel_launches", C.c_uint64), ("refined_elems", C.c_uint64),
        ("finish_runs", C.c_uint64), ("num_docs", C.c_uint64), ("ms_total", C.c_float), ("ms_h2d", C.c_float),
        ("ms_pack", C.c_float), ("ms_hist", C.c_float), ("ms_sort0", C.c_float), ("ms_refine", C.c_float),
        ("ms_finish", C.c_float), ("ms_docs", C.c_float), ("ms_remap", C.c_float), ("ms_onesweep_sum", C.c_float),
        ("ms_query", C.c_float), ("ms_hits", C.c_float), ("onesweep_elems", C.c_uint64),
        ("code_bits", C.c_uint32), ("csv_host_scan", C.c_uint32), ("ms_csv_scan", C.c_float), ("reserved1", C.c_uint32),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class SaCudaError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libsa_cuda error {code}: {msg}")
        self.code = code


def build_library(force: bool = False) -> None:
    """Compile libsa_cuda.so for sm_100a (nvcc cross-compiles without a GPU)."""
    if force:
        subprocess.run(["make", "-C", CSRC, "clean"], check=True, capture_output=True)
    r = subprocess.run(["make", "-C", CSRC, "-j8"], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("building libsa_cuda.so failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])


def build_cython_extension() -> str:
    """Cythonize + compile suffix_array_cy.pyx (the reference's host language for this layer) in-tree,
    linked against libsa_cuda.so next to it ($ORIGIN rpath).  -> path of the extension module."""
    import sysconfig
    src = os.path.join(_HERE, "suffix_array_cy.pyx")
    bdir = os.path.join(CSRC, "build")
    os.makedirs(bdir, exist_ok=True)
    c_file = os.path.join(bdir, "suffix_array_cy.c")
    out = os.path.join(_HERE, "suffix_array_cy" + sysconfig.get_config_var("EXT_SUFFIX"))
    if os.path.exists(out) and os.path.getmtime(out) >= max(os.path.getmtime(src), os.path.getmtime(LIB_PATH)):
        return out
    inc = os.path.join(os.path.dirname(_HERE), "include")
    cmds = [[sys.executable, "-m", "cython", "-3", "-I", inc, src, "-o", c_file],
            ["gcc", "-shared", "-fPIC", "-O2", "-Wno-unused-function", "-I", inc, "-I", sysconfig.get_paths()["include"],
             c_file, "-L", _HERE, "-lsa_cuda", "-Wl,-rpath,$ORIGIN", "-o", out]]
    for cmd in cmds:
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("building suffix_array_cy failed:\n" + " ".join(cmd) + "\n" + r.stdout[-3000:] + r.stderr[-3000:])
    return out


_lib = None
_vp = C.c_void_p


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
            "(the CUDA extension is required; there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    L.sa_last_error.restype = C.c_char_p
    L.sa_abi_version.restype = C.c_int
    L.sa_create.argtypes = [C.POINTER(_vp), C.c_uint32, C.c_int]
    L.sa_destroy.argtypes = [_vp]
    L.sa_destroy.restype = None
    L.sa_build_text.argtypes = [_vp, _vp, C.c_uint64, C.c_uint32]
    L.sa_build_text_device.argtypes = [_vp, _vp, C.c_uint64, C.c_uint32]
    L.sa_build_text_tail.argtypes = [_vp, _vp, C.c_uint64, C.c_uint64, C.c_uint32]
    L.sa_build_csv.argtypes = [_vp, C.c_char_p, C.c_char_p, C.c_uint32]
    L.sa_build_csv_partition.argtypes = [_vp, C.c_char_p, C.c_char_p, C.c_uint64, C.c_uint32, C.POINTER(C.c_uint64),
                                         C.POINTER(C.c_uint64)]
    L.sa_build_csv_columns.argtypes = [C.POINTER(_vp), C.POINTER(C.c_char_p), C.c_uint32, C.c_char_p, C.c_uint32]
    L.sa_size.argtypes = [_vp]
    L.sa_size.restype = C.c_uint64
    L.sa_num_docs.argtypes = [_vp]
    L.sa_num_docs.restype = C.c_uint64
    for name in ("sa_copy_sa_u32", "sa_copy_sa_i64", "sa_copy_text_positions_u32", "sa_copy_quoted_bits",
                 "sa_copy_text", "sa_copy_doc_starts", "sa_debug_copy_lcp"):
        getattr(L, name).argtypes = [_vp, _vp, C.c_uint64]
    L.sa_device_sa.argtypes = [_vp]
    L.sa_device_sa.restype = _vp
    L.sa_get_stats.argtypes = [_vp, C.POINTER(SaStats)]
    L.sa_query_ranges.argtypes = [_vp, _vp, _vp, C.c_uint64, C.c_uint32, _vp, _vp, _vp]
    L.sa_query_hits.argtypes = [_vp, _vp, _vp, C.c_uint64]
    L.sa_query_ranges_device.argtypes = [_vp, _vp, _vp, C.c_uint64, C.c_uint64, C.c_uint32, _vp, _vp]
    L.sa_hit_offsets_noalloc.argtypes = [_vp, _vp, _vp, C.c_uint64, C.c_uint32, _vp, _vp]
    L.sa_debug_set.argtypes = [_vp, C.c_char_p, C.c_int64]
    L.sa_shard_begin.argtypes = [_vp, _vp, C.c_uint64, C.c_uint32]
    L.sa_shard_begin_csv.argtypes = [_vp, C.c_char_p, C.c_char_p, C.c_uint32]
    L.sa_shard_text_size.argtypes = [_vp]
    L.sa_shard_text_size.restype = C.c_uint64
    L.sa_shard_copy_text.argtypes = [_vp, C.c_uint64, C.c_uint64, _vp]
    L.sa_shard_pack.argtypes = [_vp, C.c_uint64, C.c_uint64, _vp, _vp, _vp]
    L.sa_shard_split.argtypes = [_vp, _vp, _vp, C.c_uint64, _vp, C.c_int, _vp, _vp, _vp]
    L.sa_shard_sort.argtypes = [_vp, _vp, _vp, C.c_uint64]
    L.sa_shard_edges.argtypes = [_vp, C.c_uint32, _vp, _vp, _vp, _vp]
    L.sa_shard_finish.argtypes = [_vp, _vp, _vp, C.c_uint32, C.c_uint32, _vp, _vp, C.c_uint32, _vp]
    L.sa_adopt_sa.argtypes = [_vp, _vp, C.c_uint64]
    L.sa_adopt_sa_parts.argtypes = [_vp, _vp, C.c_uint64, _vp, C.c_int]
    L.sa_shard_count.argtypes = [_vp, _vp, C.c_uint64, _vp, C.c_int, _vp]
    L.sa_shard_recv_prepare.argtypes = [_vp, C.c_uint64, _vp, _vp]
    L.sa_shard_scatter_p2p.argtypes = [_vp, _vp, _vp, C.c_uint64, C.c_int, C.c_int, _vp, _vp]
    L.sa_shard_sort_recv.argtypes = [_vp, C.c_uint64]
    L.sa_shard_final_prepare.argtypes = [_vp, _vp]
    L.sa_shard_finish_bcast.argtypes = [_vp, _vp, _vp, C.c_uint32, C.c_uint32, _vp, _vp, C.c_uint32, C.c_int, C.c_int,
                                        _vp, C.c_uint64]
    L.sa_shard_mark_built.argtypes = [_vp]
    L.sa_save.argtypes = [_vp, C.c_char_p]
    L.sa_load.argtypes = [_vp, C.c_char_p]
    L.sa_export_reference_files.argtypes = [_vp, C.c_char_p, C.c_char_p]
    L.sa_result_create.argtypes = [_vp, C.POINTER(_vp)]
    L.sa_result_destroy.argtypes = [_vp]
    L.sa_result_destroy.restype = None
    L.sa_query_ranges_r.argtypes = [_vp, _vp, _vp, _vp, C.c_uint64, C.c_uint32, C.c_uint32, _vp, _vp, _vp]
    L.sa_query_hits_r.argtypes = [_vp, _vp, C.c_uint32, _vp, _vp, C.c_uint64, _vp]
    L.sa_query_records_size.argtypes = [_vp, _vp, C.c_uint32, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.sa_query_records.argtypes = [_vp, _vp, _vp, C.c_uint64, _vp, C.c_uint64, _vp, _vp]
    L.sa_result_stats.argtypes = [_vp, C.POINTER(C.c_float), C.POINTER(C.c_float), C.POINTER(C.c_uint64)]
    L.sa_num_rows.argtypes = [_vp]
    L.sa_num_rows.restype = C.c_uint64
    L.sa_copy_row_starts.argtypes = [_vp, _vp, C.c_uint64]
    L.sa_attach_csv_file.argtypes = [_vp, C.c_char_p]
    L.sa_comm_create_local.argtypes = [C.c_int, C.POINTER(_vp)]
    L.sa_comm_create_shm.argtypes = [C.c_char_p, C.c_int, C.c_int, C.POINTER(_vp)]
    L.sa_comm_destroy.argtypes = [_vp]
    L.sa_comm_destroy.restype = None
    L.sa_comm_rank.argtypes = [_vp]
    L.sa_comm_size.argtypes = [_vp]
    L.sa_comm_barrier.argtypes = [_vp]
    L.sa_shard_build.argtypes = [_vp, _vp, _vp, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.POINTER(C.c_float)]
    L.sa_shard_info.argtypes = [_vp, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.sa_build_text_multi.argtypes = [C.POINTER(_vp), C.c_int, _vp, C.c_uint64, C.c_uint32]
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != SA_OK:
        raise SaCudaError(rc, lib().sa_last_error().decode("utf-8", "replace"))


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(_vp)


class Index:
    """Thin RAII wrapper over an ``sa_index*`` (NumPy in, NumPy out)."""

    def __init__(self, max_suffix_length: int = 64, device: int = 0):
        self._h = _vp()
        check(lib().sa_create(C.byref(self._h), max_suffix_length, device))
        self.max_suffix_length = max_suffix_length

    def hit_offsets_noalloc(self, d_start: int, d_end: int, nq: int, k: int, d_hit_off: int, stream: int = 0) -> None:
        """sa_hit_offsets_noalloc: device pointers in, nq + 1 u64 hit offsets out, queued on `stream`
        (a cudaStream_t as int); allocates nothing, does not synchronise (include/sa_cuda.h)."""
        check(lib().sa_hit_offsets_noalloc(self._h, _vp(d_start), _vp(d_end), nq, k, _vp(d_hit_off), _vp(stream)))

    def close(self):
        r = getattr(self, "_res", None)
        if r is not None:
            r.close()
            self._res = None
        if getattr(self, "_h", None) is not None and self._h:
            lib().sa_destroy(self._h)
            self._h = None

    __del__ = close

    # ---- builds
    def build_text(self, text, lowercase: bool = False, timing: bool = False, tail: int = 0):
        """tail: the last `tail` bytes of `text` are look-ahead only (the next partition's first bytes)."""
        t = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray, memoryview)) \
            else np.ascontiguousarray(text, dtype=np.uint8)
        flags = (SA_FLAG_LOWERCASE if lowercase else 0) | (SA_FLAG_TIMING if timing else 0)
        check(lib().sa_build_text_tail(self._h, _ptr(t) if t.size else None, t.size - tail, tail, flags))

    def build_text_ptr(self, host_ptr: int, n: int, flags: int = 0):
        check(lib().sa_build_text(self._h, host_ptr, n, flags))

    def build_text_device(self, dev_ptr: int, n: int, flags: int = 0):
        check(lib().sa_build_text_device(self._h, dev_ptr, n, flags))

    def build_csv(self, path: str, column: str, timing: bool = False):
        check(lib().sa_build_csv(self._h, path.encode(), column.encode(), SA_FLAG_TIMING if timing else 0))

    def build_csv_partition(self, path: str, column: str, partition_start: int = 0, timing: bool = False):
        """One 2 GiB partition of a (big) file, engine.c:1469-1481 -> (partition_end, file_bytes)."""
        end, size = C.c_uint64(0), C.c_uint64(0)
        check(lib().sa_build_csv_partition(self._h, path.encode(), column.encode(), partition_start,
                                           SA_FLAG_TIMING if timing else 0, C.byref(end), C.byref(size)))
        return int(end.value), int(size.value)

    @staticmethod
    def build_csv_columns(indexes, path: str, columns, timing: bool = False):
        """indexes[c] <- column columns[c] of the file: one read / upload / quote analysis for all."""
        n = len(indexes)
        assert n == len(columns) and n > 0
        hs = (_vp * n)(*[ix._h for ix in indexes])
        cs = (C.c_char_p * n)(*[c.encode() for c in columns])
        check(lib().sa_build_csv_columns(hs, cs, n, path.encode(), SA_FLAG_TIMING if timing else 0))

    def save(self, path: str):
        check(lib().sa_save(self._h, path.encode()))

    def load(self, path: str):
        check(lib().sa_load(self._h, path.encode()))

    def export_reference_files(self, sa_path: str, quoted_path: str | None = None):
        check(lib().sa_export_reference_files(self._h, sa_path.encode(), quoted_path.encode() if quoted_path else None))

    def debug_set(self, key: str, value: int):
        check(lib().sa_debug_set(self._h, key.encode(), value))

    # ---- accessors
    @property
    def n(self) -> int:
        return int(lib().sa_size(self._h))

    @property
    def num_docs(self) -> int:
        return int(lib().sa_num_docs(self._h))

    def _copy(self, fn, dtype, count):
        out = np.empty(max(count, 1), dtype=dtype)
        check(fn(self._h, _ptr(out), count))
        return out[:count]

    def sa(self) -> np.ndarray:
        return self._copy(lib().sa_copy_sa_u32, np.uint32, self.n)

    def sa_i64(self) -> np.ndarray:
        return self._copy(lib().sa_copy_sa_i64, np.int64, self.n)

    def text_positions(self) -> np.ndarray:
        return self._copy(lib().sa_copy_text_positions_u32, np.uint32, self.n)

    def text(self) -> np.ndarray:
        return self._copy(lib().sa_copy_text, np.uint8, self.n)

    def lcp(self) -> np.ndarray:
        return self._copy(lib().sa_debug_copy_lcp, np.uint16, self.n)

    def doc_starts(self) -> np.ndarray:
        return self._copy(lib().sa_copy_doc_starts, np.uint32, self.num_docs)

    def quoted_bits(self) -> np.ndarray:
        nbytes = (self.n + 7) // 8
        out = np.zeros(max(nbytes, 1), dtype=np.uint8)
        check(lib().sa_copy_quoted_bits(self._h, _ptr(out), nbytes))
        return np.unpackbits(out[:nbytes], bitorder="little")[: self.n]

    def stats(self) -> dict:
        s = SaStats()
        check(lib().sa_get_stats(self._h, C.byref(s)))
        return s.as_dict()

    # ---- queries
    def query_ranges_packed(self, qbytes: np.ndarray, qoff: np.ndarray, k: int = 1000, want_offsets: bool = True):
        nq = qoff.size - 1
        start = np.empty(max(nq, 1), dtype=np.uint32)
        end = np.empty(max(nq, 1), dtype=np.uint32)
        hit_off = np.zeros(nq + 1, dtype=np.uint64) if want_offsets else None
        check(lib().sa_query_ranges(self._h, _ptr(qbytes), _ptr(qoff), nq, k, _ptr(start), _ptr(end),
                                    _ptr(hit_off) if want_offsets else None))
        return start[:nq], end[:nq], hit_off

    def query_ranges(self, queries, k: int = 1000):
        """queries: list[bytes] -> (start u32, end u32, hit_off u64[nq+1])."""
        lens = np.fromiter((len(q) for q in queries), dtype=np.uint64, count=len(queries))
        qoff = np.zeros(len(queries) + 1, dtype=np.uint64)
        np.cumsum(lens, out=qoff[1:])
        joined = b"".join(queries)
        qbytes = np.frombuffer(joined, dtype=np.uint8) if joined else np.zeros(1, np.uint8)
        return self.query_ranges_packed(qbytes, qoff, k)

    def query_hits(self, total: int):
        pos = np.empty(max(total, 1), dtype=np.uint32)
        doc = np.empty(max(total, 1), dtype=np.uint32)
        check(lib().sa_query_hits(self._h, _ptr(pos), _ptr(doc), total))
        return pos[:total], doc[:total]

    # ---- CSV row table
    @property
    def num_rows(self) -> int:
        return int(lib().sa_num_rows(self._h))

    def row_starts(self) -> np.ndarray:
        """u64[num_rows + 1]: first byte of every row of the file (row 0 = header), then the file size."""
        rows = self.num_rows
        out = np.zeros(rows + 1, dtype=np.uint64)
        check(lib().sa_copy_row_starts(self._h, _ptr(out), rows + 1))
        return out

    def attach_csv_file(self, path: str):
        check(lib().sa_attach_csv_file(self._h, path.encode()))

    def result(self) -> "Result":
        return Result(self)

    # ---- the backend surface of suffixarray_b200/_host.py
    def _own_result(self) -> "Result":
        r = getattr(self, "_res", None)
        if r is None or r._h is None:
            r = self._res = Result(self)
        return r

    def query(self, queries, k: int = 1000, file_semantics: bool = False, dedupe: bool = False):
        """-> (start, end, hit_off, hit_pos, hit_doc)"""
        r = self._own_result()
        s, e, _ = r.ranges(queries, k, SA_QUERY_FILE_SEMANTICS if file_semantics else 0)
        off, pos, doc = r.hits(dedupe)
        return s, e, off, pos, doc

    def query_records(self, queries, k: int = 1000, file_semantics: bool = False, dedupe: bool = False,
                      compat: bool = False):
        """-> (hit_off, rec_off, bytes u8, record ids)"""
        r = self._own_result()
        r.ranges(queries, k, SA_QUERY_FILE_SEMANTICS if file_semantics else 0)
        return r.records(dedupe, compat)


SHARD_PHASES = ("text", "prepare", "pack+hist", "split", "exchange", "sort+refine", "edges", "finish+replicate")


class Comm:
    """Node-local communicator of the native multi-GPU driver (``sa_comm``): one per rank (process)."""

    def __init__(self, name: str, rank: int, nranks: int):
        self._h = _vp()
        check(lib().sa_comm_create_shm(name.encode(), rank, nranks, C.byref(self._h)))
        self.rank, self.nranks = rank, nranks

    def barrier(self):
        check(lib().sa_comm_barrier(self._h))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            lib().sa_comm_destroy(self._h)
            self._h = None

    __del__ = close


def shard_build(index: "Index", comm: Comm, slice_ptr: int, slice_off: int, slice_len: int, n: int, flags: int = 0):
    """One rank of the native sample-sort build (sa_shard_build).  slice_ptr: host address of this rank's
    slice of the text (device address with SA_FLAG_DEVICE_TEXT).  -> dict of per-phase host wall ms."""
    ph = (C.c_float * 16)()
    check(lib().sa_shard_build(index._h, comm._h, slice_ptr, slice_off, slice_len, n, flags, ph))
    return {name: float(ph[i]) for i, name in enumerate(SHARD_PHASES)}


def shard_info(index: "Index"):
    a, b = C.c_uint64(0), C.c_uint64(0)
    check(lib().sa_shard_info(index._h, C.byref(a), C.byref(b)))
    return int(a.value), int(b.value)


def build_text_multi(indexes, text, lowercase: bool = False, timing: bool = False, sharded_sa: bool = False):
    """ONE index over the GPUs of `indexes` (handles of one process, a thread per handle): every handle
    ends up with the text and the complete array."""
    t = np.frombuffer(text, dtype=np.uint8) if isinstance(text, (bytes, bytearray, memoryview)) \
        else np.ascontiguousarray(text, dtype=np.uint8)
    n = len(indexes)
    hs = (_vp * n)(*[ix._h for ix in indexes])
    flags = (SA_FLAG_LOWERCASE if lowercase else 0) | (SA_FLAG_TIMING if timing else 0) | (SA_FLAG_SHARDED_SA if sharded_sa else 0)
    check(lib().sa_build_text_multi(hs, n, _ptr(t) if t.size else None, t.size, flags))


def _pack(queries):
    lens = np.fromiter((len(q) for q in queries), dtype=np.uint64, count=len(queries))
    qoff = np.zeros(len(queries) + 1, dtype=np.uint64)
    np.cumsum(lens, out=qoff[1:])
    joined = b"".join(queries)
    qbytes = np.frombuffer(joined, dtype=np.uint8) if joined else np.zeros(1, np.uint8)
    return qbytes, qoff


class Result:
    """Per-call query state (``sa_result``): one batch at a time per Result, any number of Results
    (threads) per built Index."""

    def __init__(self, index: Index):
        self._index = index
        self._h = _vp()
        check(lib().sa_result_create(index._h, C.byref(self._h)))
        self.nq = 0
        self.total_hits = 0

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            lib().sa_result_destroy(self._h)
            self._h = None

    __del__ = close

    def ranges_packed(self, qbytes: np.ndarray, qoff: np.ndarray, k: int = 1000, flags: int = 0):
        nq = qoff.size - 1
        start = np.empty(max(nq, 1), dtype=np.uint32)
        end = np.empty(max(nq, 1), dtype=np.uint32)
        hit_off = np.zeros(nq + 1, dtype=np.uint64)
        check(lib().sa_query_ranges_r(self._index._h, self._h, _ptr(qbytes), _ptr(qoff), nq, k, flags, _ptr(start),
                                      _ptr(end), _ptr(hit_off)))
        self.nq = nq
        self.total_hits = int(hit_off[nq])
        return start[:nq], end[:nq], hit_off

    def ranges(self, queries, k: int = 1000, flags: int = 0):
        qbytes, qoff = _pack(queries)
        return self.ranges_packed(qbytes, qoff, k, flags)

    def hits(self, dedupe: bool = False):
        """-> (hit_off u64[nq+1], hit_pos u32, hit_doc u32) of the last ranges() call."""
        total = self.total_hits
        pos = np.empty(max(total, 1), dtype=np.uint32)
        doc = np.empty(max(total, 1), dtype=np.uint32)
        off = np.zeros(self.nq + 1, dtype=np.uint64)
        check(lib().sa_query_hits_r(self._index._h, self._h, SA_QUERY_DEDUPE if dedupe else 0, _ptr(pos), _ptr(doc), total,
                                    _ptr(off)))
        out = int(off[self.nq]) if self.nq else 0
        return off, pos[:out], doc[:out]

    def records(self, dedupe: bool = False, compat: bool = False):
        """-> (hit_off u64[nq+1], rec_off u64[records+1], bytes u8, record ids u32): every hit's record
        (document / line of the CSV file) gathered on the device into one packed buffer."""
        flags = (SA_QUERY_DEDUPE if dedupe else 0) | (SA_RECORDS_REFERENCE_COMPAT if compat else 0)
        num, nbytes = C.c_uint64(0), C.c_uint64(0)
        check(lib().sa_query_records_size(self._index._h, self._h, flags, C.byref(num), C.byref(nbytes)))
        num, nbytes = int(num.value), int(nbytes.value)
        out = np.empty(max(nbytes, 1), dtype=np.uint8)
        rec_off = np.zeros(num + 1, dtype=np.uint64)
        hit_off = np.zeros(self.nq + 1, dtype=np.uint64)
        doc = np.empty(max(num, 1), dtype=np.uint32)
        check(lib().sa_query_records(self._index._h, self._h, _ptr(out), nbytes, _ptr(rec_off), num, _ptr(hit_off), _ptr(doc)))
        return hit_off, rec_off, out[:nbytes], doc[:num]

    def stats(self) -> dict:
        a, b, n = C.c_float(0), C.c_float(0), C.c_uint64(0)
        check(lib().sa_result_stats(self._h, C.byref(a), C.byref(b), C.byref(n)))
        return {"ms_query": a.value, "ms_hits": b.value, "kernel_launches": int(n.value)}
