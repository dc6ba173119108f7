import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "datagen")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")

# test infrastructure: if the CUDA library has not been built yet (fresh checkout) and nvcc is here, build it; the
# package itself never builds or falls back -- it refuses to load without lib/libwgbv.so
if not os.path.exists(os.path.join(ROOT, "webgraph-big_b200", "lib", "libwgbv.so")):
    import shutil
    if shutil.which("nvcc") or os.path.exists("/usr/local/cuda/bin/nvcc"):
        sys.path.insert(0, ROOT)
        import __graft_entry__
        __graft_entry__.build()


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_cnr_golden():
    """(deg int64[n], rowptr int64[n+1], col int64[m]) of the reference's cnr-2000 ASCII golden (see tests/golden/make_golden.py)."""
    z = np.load(os.path.join(GOLDEN, "cnr-2000.golden.npz"))
    deg = z["deg"].astype(np.int64)
    cd = z["col_delta"].astype(np.int64)
    rowptr = np.zeros(deg.size + 1, dtype=np.int64)
    np.cumsum(deg, out=rowptr[1:])
    c = np.cumsum(cd)
    starts = rowptr[:-1][deg > 0]
    sub = np.zeros(deg.size, dtype=np.int64)
    sub[deg > 0] = c[starts] - cd[starts]
    col = c - sub[np.repeat(np.arange(deg.size), deg)]
    return deg, rowptr, col


@pytest.fixture(scope="session")
def cnr_golden():
    return load_cnr_golden()


@pytest.fixture(scope="session")
def cnr_basename():
    return os.path.join(GOLDEN, "cnr-2000")
