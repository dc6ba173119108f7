"""CPU-side checks of the C ABI: the library loads, exports every symbol include/wgbv.h declares, and the
host-only entry points (properties parsing, argument checks) behave like the reference.  No GPU needed."""
import ctypes as C
import os
import re

import pytest

import webgraph_big_b200 as wg
from webgraph_big_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "wgbv.h")).read()
    return sorted(set(re.findall(r"WGBV_API[^;(]*?\b(wgbv_\w+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    names = declared_symbols()
    assert len(names) >= 25
    L = C.CDLL(_lib.LIB_PATH)
    for n in names:
        assert hasattr(L, n), n
    assert sorted(n for n, _, _ in _lib.SYMBOLS) == names  # the ctypes table covers the whole header


def test_parse_properties_matches_reference_checks(cnr_basename):
    p = wg.parse_properties(open(cnr_basename + ".properties").read())
    assert (p.nodes, p.arcs, p.window_size, p.max_ref_count, p.min_interval_length, p.zeta_k) == (325557, 3216152, 7, 3, 3, 3)
    assert (p.outdegree_coding, p.block_coding, p.residual_coding, p.reference_coding, p.block_count_coding, p.offset_coding) == (2, 2, 6, 5, 2, 2)
    base = "graphclass=it.unimi.dsi.webgraph.BVGraph\nversion=0\nnodes=1\narcs=0\nwindowsize=7\nmaxrefcount=3\nminintervallength=4\n"
    q = wg.parse_properties(base + "compressionflags=RESIDUALS_DELTA | OFFSETS_DELTA|BLOCK_COUNT_UNARY\n")
    assert (q.residual_coding, q.offset_coding, q.block_count_coding, q.zeta_k) == (1, 1, 5, 3)
    with pytest.raises(wg.IOException):  # BVG:1491
        wg.parse_properties(base.replace("BVGraph", "EFGraph"))
    with pytest.raises(wg.IOException):  # BVG:1496
        wg.parse_properties(base.replace("version=0\n", ""))
    with pytest.raises(wg.IOException):  # BVG:1497
        wg.parse_properties(base.replace("version=0", "version=1"))
    with pytest.raises(wg.IOException):  # BVG:1327
        wg.parse_properties(base + "compressionflags=RESIDUALS_FOO\n")


def test_missing_files_raise_io(tmp_path):
    with pytest.raises(wg.IOException):
        wg.GpuBVGraph.load(str(tmp_path / "nothing"))


def test_null_arguments_are_rejected():
    L = _lib.lib()
    assert L.wgbv_stream_create(None, None, None) == _lib.EARG
    assert L.wgbv_decode_csr(None, 0, 0, None, None, 0) == _lib.EARG
    assert b"null" in L.wgbv_last_error()
