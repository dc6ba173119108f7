"""The C++ host mirror compiles against the C ABI (syntax + link check against libwgbv.so; no GPU needed) and the Java
shim source names only entry points that include/wgbv.h declares."""
import os
import re
import subprocess
import sys

import pytest

from webgraph_big_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cpp_mirror_compiles_and_links(tmp_path):
    src = tmp_path / "t.cpp"
    src.write_text('#include "wgbv.hpp"\n'
                   'int main(int argc, char **argv) {\n'
                   '  try { auto g = wgbv::Graph::load(argc > 1 ? argv[1] : "/nonexistent/graph"); return (int)g->numNodes(); }\n'
                   '  catch (const wgbv::IOException &) { return 42; }\n'
                   '  catch (const wgbv::CudaError &) { return 43; }\n'
                   '}\n')
    exe = tmp_path / "t"
    libdir = os.path.dirname(_lib.LIB_PATH)
    subprocess.check_call(["g++", "-std=c++17", "-Wall", "-Werror", "-I", os.path.join(ROOT, "webgraph-big_b200", "host"), str(src),
                           "-L", libdir, "-lwgbv", f"-Wl,-rpath,{libdir}", "-o", str(exe)])
    assert subprocess.call([str(exe)]) == 42  # missing files are the reference's IOException


def test_java_shim_only_names_declared_entry_points():
    text = open(os.path.join(ROOT, "webgraph-big_b200", "java", "GpuBVGraph.java")).read()
    used = set(re.findall(r'h\("(wgbv_\w+)"', text))
    declared = {n for n, _, _ in _lib.SYMBOLS}
    assert used and used <= declared


def test_write_offsets_reproduces_the_reference_fixture():
    """wgbv_write_offsets = BVGraph.writeOffsets (BVG:2595-2609): from cnr-2000.properties and cnr-2000.graph alone it
    writes the reference's own cnr-2000.offsets byte for byte; for every coding the format allows it writes the
    offsets stream the generator wrote; a truncated graph is refused.  Host code (no GPU involved)."""
    import numpy as np
    import webgraph_big_b200 as wg
    import bvgen as bg
    base = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "cnr-2000")
    props = open(base + ".properties").read()
    graph = np.fromfile(base + ".graph", dtype=np.uint8)
    want = np.fromfile(base + ".offsets", dtype=np.uint8)
    assert np.array_equal(wg.write_offsets(props, graph), want)
    combos = [dict(residual=bg.ZETA, zetak=1), dict(residual=bg.ZETA, zetak=7), dict(residual=bg.GAMMA), dict(residual=bg.DELTA),
              dict(residual=bg.NIBBLE), dict(residual=bg.GOLOMB, zetak=5), dict(outdegree=bg.DELTA, block=bg.DELTA, blockcount=bg.DELTA),
              dict(blockcount=bg.UNARY, reference=bg.GAMMA), dict(reference=bg.DELTA, offset=bg.DELTA),
              dict(window=0), dict(window=1, maxref=1, minint=0), dict(window=12, maxref=8, minint=2)]
    for i, kw in enumerate(combos):
        e = bg.copy_model(6000, seed=50 + i, mean_degree=22, params=bg.make_params(**kw), nsplits=2)
        assert np.array_equal(wg.write_offsets(e.properties_text(), e.graph), np.asarray(e.offsets_stream)), kw
    e = bg.copy_model(3000, seed=2, mean_degree=10, nsplits=1)
    with pytest.raises(wg.IOException):
        wg.write_offsets(e.properties_text(), e.graph[: e.graph.size // 2])
