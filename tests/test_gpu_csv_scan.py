"""Device CSV column scanner (csv_scan.cu) vs the oracle's restatement of the reference's byte
state machine (oracle/sa_oracle.c sao_csv_scan <-> engine.c:522-605): column text, file offsets
and quoted flags, for regular files through the device scan, and for irregular files through the
automatic host-scan route."""
import os
import tempfile

import numpy as np
import pytest

from oracle import pyoracle as O
from suffixarray_b200 import workloads as W

pytestmark = pytest.mark.gpu

L = 32


def expected(data: bytes, col: int, ncols: int):
    text, mp, qd = O.oracle_csv_scan(data, col, ncols)
    sa = O.oracle_build(text, L)
    sa_file, q_rank = O.oracle_csv_remap(sa, mp, qd)
    return text, sa, sa_file, q_rank


def build(data: bytes, column: str, scanner: int):
    from suffixarray_b200 import _lib as S
    with tempfile.NamedTemporaryFile(suffix=".csv", delete=False) as f:
        f.write(data)
        path = f.name
    idx = S.Index(L)
    try:
        idx.debug_set("csv_scanner", scanner)
        idx.build_csv(path, column)
    except Exception:
        idx.close()
        raise
    finally:
        os.unlink(path)
    return idx


def check(data: bytes, header: list[str], scanner: int, want_host: int | None = None):
    for col, name in enumerate(header):
        text, sa, sa_file, q_rank = expected(data, col, len(header))
        idx = build(data, name, scanner)
        try:
            assert idx.n == text.size, (name, idx.n, text.size)
            assert np.array_equal(idx.text(), text), name
            assert np.array_equal(idx.text_positions(), sa), name
            assert np.array_equal(idx.sa(), sa_file), name           # = map[SA]: checks every file offset
            assert np.array_equal(idx.quoted_bits(), q_rank), name
            if want_host is not None:
                assert idx.stats()["csv_host_scan"] == want_host
        finally:
            idx.close()


REGULAR = {
    "plain": b"a,b,c\nx1,y1,z1\nx2,y2,z2\n",
    "no_trailing_newline": b"a,b,c\nx1,y1,z1\nx2,y2,z2",
    "no_trailing_newline_quoted": b'a,b,c\nx1,y1,"z,1"\nx2,y2,"z2"',
    "ends_with_comma": b"a,b,c\nx1,y1,z1\nx2,y2,",
    "empty_cells": b"a,b,c\n,,\nx,,z\n,y,\n",
    "quoted": b'a,b,c\n"hello, world",y1,"z""1"\n"",y2,"""q"""\n"multi\nline",y3,z3\n',
    "upper_and_crlf": b'a,b,c\r\nHeLLo,"WORLD",Z\r\n"Fo""O",Bar,baz\r\n',
    "ragged_rows": b"a,b,c\nx1,y1\nx2,y2,z2,w2\nx3\n",
    "one_column": b"a\nfoo\nbar\n\nbaz\n",
    "only_header": b"a,b,c\n",
    "only_header_no_newline": b"a,b,c",
    "escaped_pairs_at_edges": b'a,b\n"""x""","y"""\n"""""",""""\n',
    "high_bytes": b"a,b\n\xff\xfe,caf\xc3\xa9\n\x80,\x7f\n",
}


@pytest.mark.parametrize("name", sorted(REGULAR))
@pytest.mark.parametrize("scanner", [2, 1])
def test_regular_files(name, scanner):
    data = REGULAR[name]
    header = data.split(b"\n")[0].rstrip(b"\r").decode().split(",")
    check(data, header, scanner, want_host=1 if scanner == 1 else 0)


IRREGULAR = {
    "quote_inside_unquoted_cell": b'a,b,c\nab"c,d"e,f\nx,y,z\n',
    "text_after_closing_quote": b'a,b,c\n"ab"c,d,e\nx,y,z\n',
    "unterminated_quote": b'a,b,c\nx,"y,z\nq,r,s\n',
    "nul_byte": b"a,b,c\nx,y\x00y,z\nq,r,s\n",
    "lone_quote_at_cell_end": b'a,b,c\nx,y",z\nq,r,s\n',
    # the reference only closes a quoted cell in front of , \n or the end of the file (engine.c:570):
    # a quoted last cell of a CRLF row runs on into the next row
    "crlf_after_closing_quote": b'a,b,c\r\nFoo,Bar,"Ba""Z"\r\nx,y,z\r\n',
}


@pytest.mark.parametrize("name", sorted(IRREGULAR))
def test_irregular_files_take_the_host_scan(name):
    data = IRREGULAR[name]
    check(data, ["a", "b", "c"], 0, want_host=1)
    with pytest.raises(Exception):
        build(data, "b", 2)


@pytest.mark.parametrize("size", [4095, 4096, 4097, 4112, 8192, 12288 - 1, 65536, 65536 + 15])
def test_file_sizes_around_tile_edges(size):
    """the device scan works on 4096-byte tiles of 16-byte groups: exact multiples, one off,
    with and without a trailing newline."""
    base = W.csv_file_bytes(max(8, size // 20), seed=size, quoted_fraction=0.3)
    data = base[:size]
    if data.count(b'"') % 2:                      # keep the quoting regular after the cut
        cut = data.rfind(b'"')
        data = data[:cut] + b"x" + data[cut + 1:]
    header = ["id", "name", "city", "score"]
    try:
        check(data, header, 2, want_host=0)
    except Exception as exc:                      # the cut may still have produced an irregular cell
        if "not a regular CSV" not in str(exc):
            raise
        check(data, header, 0, want_host=1)


def test_generated_csv_both_scanners_agree():
    data = W.csv_file_bytes(20_000, seed=77, quoted_fraction=0.25)
    header = ["id", "name", "city", "score"]
    check(data, header, 2, want_host=0)
    a = build(data, "name", 2)
    b = build(data, "name", 1)
    try:
        assert np.array_equal(a.sa(), b.sa()) and np.array_equal(a.text(), b.text())
        assert np.array_equal(a.quoted_bits(), b.quoted_bits())
    finally:
        a.close()
        b.close()
