"""The C++ host mirror compiles against the C ABI (syntax + link check against libwgbv.so; no GPU needed) and the Java
shim source names only entry points that include/wgbv.h declares."""
import os
import re
import subprocess

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
