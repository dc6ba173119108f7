"""Builds libwgbv.so (the CUDA decode engine + its C ABI) in-tree for sm_100a with nvcc.

The library is plain CUDA runtime + C++17: no torch headers, no Python.  It is written to
``webgraph-big_b200/lib/libwgbv.so`` so that it travels with the repository snapshot to the GPU box.
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libwgbv.so")
SOURCES = ["wgbv_api.cu", "bv_kernels.cu", "bv_fused.cu", "bv_encode.cu", "bv_transpose.cu", "bv_offsets.cu", "bv_labels.cu", "bv_noalloc.cu", "bv_loader.cpp"]
HEADERS = ["bv_device.cuh", "bv_rows.cuh", "bv_kernels.cuh", "bv_block.cuh", "bv_loader.h", os.path.join("..", "..", "include", "wgbv.h")]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-Wall,-Wextra,-Wno-parentheses,-fvisibility=hidden", "-shared", "-cudart", "static"]


def _nvcc():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def needs_build():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False, out=None, extra_flags=()):
    """out / extra_flags: an experimental variant under another name (A/B runs through WGBV_LIBRARY), e.g.
    ``python build.py --variant name -DSOME_SWITCH=1`` writes lib/libwgbv_name.so."""
    if out is None and not force and not needs_build():
        return LIB
    os.makedirs(LIBDIR, exist_ok=True)
    cmd = [_nvcc()] + NVCC_FLAGS + list(extra_flags) + (["-Xptxas", "-v"] if verbose else []) + \
          [os.path.join(CSRC, f) for f in SOURCES] + ["-o", out or LIB]
    res = subprocess.run(cmd, cwd=CSRC, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or res.returncode:
        sys.stderr.write(res.stdout)
    if res.returncode:
        raise RuntimeError("nvcc failed building libwgbv.so")
    return out or LIB


if __name__ == "__main__":
    if "--variant" in sys.argv:
        i = sys.argv.index("--variant")
        print(build(verbose="-v" in sys.argv, out=os.path.join(LIBDIR, "libwgbv_%s.so" % sys.argv[i + 1]),
                    extra_flags=[a for a in sys.argv[i + 2:] if a.startswith("-D")]))
    else:
        print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
