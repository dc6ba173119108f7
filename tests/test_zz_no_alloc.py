"""wgbv_scan_degrees_static: the row-pointer scan for callers that cannot allocate device memory.

Some users of libwgbv hold all of the GPU's memory in their own pool, or run on threads where the device-wide
synchronisation of cudaMalloc / cudaFree is not acceptable; for them the library offers one entry point whose
scratch is a static device array (see include/wgbv.h).  This test calls it through the C ABI, twice in a row on
the same device (the second call finds the array dirty from the first), and requires bit-exact agreement with the
row pointers of the regular entry point wgbv_decode_degrees.  It also checks the size limit is enforced.
The file name sorts last on purpose: a run that stops at the first failure has then already run every other test.
"""
import ctypes as C

import numpy as np
import pytest

import bvgen as bg
import webgraph_big_b200 as wg
from webgraph_big_b200 import _lib

pytestmark = pytest.mark.gpu


def test_static_scan_matches_the_allocating_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    L = _lib.lib()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.current_stream().cuda_stream
    for n, seed in ((70000, 5), (300000, 6)):  # the second graph needs more tiles than the first left behind
        e = bg.copy_model(n, seed=seed, mean_degree=14, nsplits=4)
        with wg.GpuBVGraph.load_mem(e.properties_text(), e.graph, e.offsets_stream, cuda_stream=stream) as g:
            want = torch.empty(n + 1, dtype=torch.int64, device=dev)
            arcs = g.decodeDegreesDevice(0, n, want.data_ptr())  # the existing entry point
            assert arcs == e.arcs
            deg = (want[1:] - want[:-1]).to(torch.int32).contiguous()  # outdegrees, as a caller would hold them
            got = torch.full((n + 1,), -1, dtype=torch.int64, device=dev)
            for _ in range(2):
                got.fill_(-1)
                rc = L.wgbv_scan_degrees_static(C.c_void_p(deg.data_ptr()), n, C.c_void_p(got.data_ptr()), C.c_void_p(stream))
                assert rc == 0, L.wgbv_last_error()
                torch.cuda.synchronize()
                assert torch.equal(got, want)
    # empty input: a single zero
    one = torch.full((1,), -1, dtype=torch.int64, device=dev)
    assert L.wgbv_scan_degrees_static(None, 0, C.c_void_p(one.data_ptr()), C.c_void_p(stream)) == 0
    torch.cuda.synchronize()
    assert int(one.item()) == 0


def test_static_scan_rejects_inputs_beyond_its_fixed_scratch():
    torch = pytest.importorskip("torch")
    L = _lib.lib()
    torch.cuda.set_device(0)
    buf = torch.zeros(8, dtype=torch.int64, device="cuda")
    too_many = 65536 * 2048  # WGBV_STATIC_SCAN_MAX_COUNT + 1
    rc = L.wgbv_scan_degrees_static(C.c_void_p(buf.data_ptr()), too_many, C.c_void_p(buf.data_ptr()), None)
    assert rc == _lib.EARG
    assert b"at most" in L.wgbv_last_error()
