"""The C-ABI library loads on a machine without a GPU and exports every symbol that
include/sa_cuda.h declares; with no device every entry point fails loudly (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

from conftest import ROOT, _cuda_available
from suffixarray_b200 import _lib as S

HEADER = os.path.join(ROOT, "include", "sa_cuda.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sa_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_are_exported():
    lib = ctypes.CDLL(S.LIB_PATH)
    names = declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in sa_cuda.h but not exported"
    assert sorted(S.EXPORTS) == names


def test_abi_version():
    assert S.lib().sa_abi_version() == 2


def test_stats_struct_matches_header():
    src = open(HEADER).read()
    body = re.search(r"typedef struct sa_stats \{(.*?)\} sa_stats;", src, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    fields = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        typ, names = decl.split(None, 1)
        fields += [(n.strip(), typ) for n in names.split(",")]
    ctype = {"uint64_t": ctypes.c_uint64, "uint32_t": ctypes.c_uint32, "float": ctypes.c_float}
    assert [(n, ctype[t]) for n, t in fields] == list(S.SaStats._fields_)


@pytest.mark.skipif(_cuda_available(), reason="checks the no-GPU failure mode")
def test_fails_loudly_without_gpu():
    with pytest.raises(S.SaCudaError) as ei:
        S.Index(32)
    assert ei.value.code == -3 and "no CPU fallback" in str(ei.value)


def test_invalid_arguments_do_not_crash():
    L = S.lib()
    assert L.sa_create(None, 32, 0) == -1
    h = ctypes.c_void_p()
    assert L.sa_create(ctypes.byref(h), 5000, 0) == -5
    assert L.sa_size(None) == 0
    assert L.sa_num_docs(None) == 0
    L.sa_destroy(None)
    out = np.zeros(4, np.uint32)
    assert L.sa_copy_sa_u32(None, out.ctypes.data_as(ctypes.c_void_p), 4) == -1


def test_multi_column_entry_validates_before_touching_the_device():
    L = S.lib()
    vp, cp = ctypes.c_void_p, ctypes.c_char_p
    assert L.sa_build_csv_columns(None, None, 1, b"x.csv", 0) == -1
    hs = (vp * 1)(None)
    cs = (cp * 1)(b"name")
    assert L.sa_build_csv_columns(hs, cs, 0, b"x.csv", 0) == -1          # no columns
    assert L.sa_build_csv_columns(hs, cs, 65, b"x.csv", 0) == -5         # more than 64 per call
    assert L.sa_build_csv_columns(hs, cs, 1, b"x.csv", 0) == -1          # NULL handle in the list
    assert b"NULL entry" in L.sa_last_error()
    assert L.sa_build_csv(None, b"x.csv", b"name", 0) == -1
    assert L.sa_shard_begin_csv(None, b"x.csv", b"name", 0) == -1
    assert L.sa_shard_text_size(None) == 0


def test_cython_host_side_builds_and_fails_loudly_without_gpu():
    """suffix_array_cy.pyx (the reference's Cython class over sa_cuda.h) cythonizes, compiles and links
    against the in-tree libsa_cuda.so; it exposes the reference's method names; without a device the
    constructor raises (no CPU fallback)."""
    path = S.build_cython_extension()
    assert os.path.exists(path)
    from suffixarray_b200 import suffix_array_cy as cy
    for name in ("construct_truncated_suffix_array_documents", "construct_truncated_suffix_array_from_csv",
                 "query_records", "query_records_batch", "save", "load", "suffix_array"):
        assert hasattr(cy.SuffixArray, name), name
    if not _cuda_available():
        with pytest.raises(cy.SaCudaError) as ei:
            cy.SuffixArray(documents=["a", "b"], max_suffix_length=8)
        assert ei.value.code == -3 and "no CPU fallback" in str(ei.value)


def test_library_is_usable_from_plain_c(tmp_path):
    """tests/c_abi_smoke.c: compiled with gcc against include/sa_cuda.h and the in-tree library -- no
    Python or torch types anywhere in the boundary.  On a machine without a GPU it must report the loud
    failure of sa_create; with one it builds and queries "banana\\nbandana" and checks the known answers."""
    import subprocess
    exe = str(tmp_path / "c_abi_smoke")
    libdir = os.path.dirname(S.LIB_PATH)
    r = subprocess.run(["gcc", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "c_abi_smoke.c"),
                        "-L", libdir, "-lsa_cuda", f"-Wl,-rpath,{libdir}", "-o", exe], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, (r.stdout, r.stderr)
    assert r.stdout.startswith("ok" if _cuda_available() else "nogpu"), r.stdout


def test_newline_tile_arithmetic_at_the_size_limit():
    """the tile count of the newline scan must not wrap for n close to 2^32 (the sharded build
    accepts n up to 2^32 - 1026).  Host arithmetic only: `sa_newline_tiles` (query.cu) through its C++ symbol."""
    L = S.lib()
    try:
        f = getattr(L, "_Z16sa_newline_tilesj")
    except AttributeError:
        pytest.skip("sa_newline_tiles is not an exported symbol in this build")
    f.argtypes = [ctypes.c_uint32]
    f.restype = ctypes.c_uint32
    tile = 4096
    for n in (0, 1, tile - 1, tile, tile + 1, (1 << 31) - 1, 1 << 31, (1 << 32) - 4096, (1 << 32) - 4095,
              (1 << 32) - 1026, (1 << 32) - 1):
        assert f(n) == (n + tile - 1) // tile, n
