"""webgraph-big_b200 -- a B200 (sm_100a) BVGraph decompression engine behind webgraph-big's ImmutableGraph API.

This module is the host-side mirror of the reference's plugin interface for the BVGraph decode path, written
over the C ABI of libwgbv (include/wgbv.h): same method names, argument meaning and error behaviour as
``it.unimi.dsi.big.webgraph.ImmutableGraph`` / ``NodeIterator`` / ``LazyLongIterator`` (reference files
ImmutableGraph.java:201-447, NodeIterator.java:34-133, LazyLongIterator.java:28-43, BVGraph.java:821-1265), plus
the bulk call ``decodeToCSR``.  Every decode runs in CUDA; nothing here decodes on the CPU.

The directory name contains a hyphen: import it as ``import webgraph_big_b200`` (alias module at the repository
root) or through ``importlib.import_module("webgraph-big_b200")``.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import (CudaError, IllegalArgumentException, IllegalStateException, IOException,  # noqa: F401
                   NoSuchElementException, Props, UnsupportedOperationException)

__all__ = ["GpuBVGraph", "NodeIterator", "LazyLongIterator", "Props", "parse_properties", "IllegalArgumentException",
           "IllegalStateException", "IOException", "UnsupportedOperationException", "NoSuchElementException", "CudaError"]

GRAPH_EXTENSION, OFFSETS_EXTENSION, PROPERTIES_EXTENSION = ".graph", ".offsets", ".properties"


def write_offsets(props, graph_bytes):
    """BVGraph.writeOffsets (BVG:2595-2609): the bytes of the .offsets file of a graph given its properties (a Props or
    the text of the .properties file) and its .graph bytes.  Host code, one sequential pass over the records."""
    if isinstance(props, str):
        props = parse_properties(props)
    g = np.ascontiguousarray(np.frombuffer(graph_bytes, dtype=np.uint8)) if not isinstance(graph_bytes, np.ndarray) else np.ascontiguousarray(graph_bytes)
    size = C.c_uint64(0)
    f = _lib.lib().wgbv_write_offsets
    _lib.check(f(C.byref(props), g.ctypes.data if g.size else None, g.size, None, 0, C.byref(size)))
    out = np.zeros(max(size.value, 1), dtype=np.uint8)
    _lib.check(f(C.byref(props), g.ctypes.data if g.size else None, g.size, out.ctypes.data, out.size, C.byref(size)))
    return out[: size.value]


def parse_properties(text):
    """The .properties checks of BVGraph.loadInternal (BVG:1482-1503) -> Props."""
    p = Props()
    _lib.check(_lib.lib().wgbv_parse_properties(text.encode("utf-8"), C.byref(p)))
    return p


class LazyLongIterator:
    """The reference's lazy iterator protocol (LazyLongIterator.java:28-43): nextLong() returns -1 when exhausted."""

    def __init__(self, values):
        self._v = values
        self._i = 0

    def nextLong(self):
        if self._i >= len(self._v):
            return -1
        r = int(self._v[self._i])
        self._i += 1
        return r

    def skip(self, n):
        k = min(int(n), len(self._v) - self._i)
        self._i += k
        return k

    def __iter__(self):
        while True:
            r = self.nextLong()
            if r == -1:
                return
            yield r


class _Stream:
    """One wgbv_stream: the mutable state behind one flyweight copy (IG:187-197)."""

    def __init__(self, graph_handle, cuda_stream=None):
        self.h = C.c_void_p()
        _lib.check(_lib.lib().wgbv_stream_create(graph_handle, C.c_void_p(cuda_stream or 0), C.byref(self.h)))

    def close(self):
        if self.h:
            _lib.lib().wgbv_stream_destroy(self.h)
            self.h = C.c_void_p()


class _Shared:
    """The immutable part shared by a graph and all its copies (bit stream + Elias-Fano offsets on the device)."""

    def __init__(self, handle, basename):
        self.handle = handle
        self.basename = basename
        self.refs = 0

    def release(self):
        self.refs -= 1
        if self.refs == 0 and self.handle:
            _lib.lib().wgbv_close(self.handle)
            self.handle = C.c_void_p()


class GpuBVGraph:
    """A BVGraph whose successor lists are decoded on a B200.

    A sibling of ``it.unimi.dsi.big.webgraph.BVGraph`` (it accepts the same files, ``graphclass`` naming BVGraph,
    standard or big), selected explicitly like ``-g GpuBVGraph`` would in the reference's tools.
    """

    CHUNK_NODES = 1 << 16  # nodes decoded per bulk call behind nodeIterator()

    def __init__(self, shared, cuda_stream=None):
        self._sh = shared
        shared.refs += 1
        self._st = _Stream(shared.handle, cuda_stream)
        L = _lib.lib()
        self._n = L.wgbv_num_nodes(shared.handle)
        self._m = L.wgbv_num_arcs(shared.handle)
        lo, hi = C.c_int64(), C.c_int64()
        _lib.check(L.wgbv_shard_range(shared.handle, C.byref(lo), C.byref(hi)))
        self._lo, self._hi = lo.value, hi.value

    # -------------------------------------------------------------------------------------------- loading
    @classmethod
    def load(cls, basename, device=0, shard_index=0, shard_count=1, cuda_stream=None):
        """BVGraph.load(basename) (BVG:1345-1383): reads basename.properties/.graph/.offsets onto the device."""
        h = C.c_void_p()
        _lib.check(_lib.lib().wgbv_open(str(basename).encode("utf-8"), device, shard_index, shard_count, C.byref(h)))
        return cls(_Shared(h, str(basename)), cuda_stream)

    # the reference's other loaders all need the whole stream in device memory here
    loadMapped = load
    loadOffline = load

    @classmethod
    def load_mem(cls, props, graph_bytes, offsets_bytes, device=0, shard_index=0, shard_count=1, cuda_stream=None):
        """Same from memory: ``props`` is a Props or the text of a .properties file.  offsets_bytes None: the graph came
        without its .offsets and the records are found by the sequential pass of BVGraph.writeOffsets (see write_offsets)."""
        if isinstance(props, str):
            props = parse_properties(props)
        g = np.ascontiguousarray(np.frombuffer(graph_bytes, dtype=np.uint8)) if not isinstance(graph_bytes, np.ndarray) else np.ascontiguousarray(graph_bytes)
        if offsets_bytes is None:
            o = None
        else:
            o = np.ascontiguousarray(np.frombuffer(offsets_bytes, dtype=np.uint8)) if not isinstance(offsets_bytes, np.ndarray) else np.ascontiguousarray(offsets_bytes)
        h = C.c_void_p()
        _lib.check(_lib.lib().wgbv_open_mem(C.byref(props), g.ctypes.data, g.size, o.ctypes.data if o is not None else None,
                                            o.size if o is not None else 0, device, shard_index, shard_count, C.byref(h)))
        return cls(_Shared(h, None), cuda_stream)

    @classmethod
    def load_slice(cls, props, first_node, node_count, bits, offsets_bytes, halo_nodes=0, device=0, cuda_stream=None):
        """Opens a pre-cut slice of a larger graph (one shard per process): see wgbv_open_slice in include/wgbv.h."""
        if isinstance(props, str):
            props = parse_properties(props)
        b = np.ascontiguousarray(bits)
        o = np.ascontiguousarray(offsets_bytes)
        h = C.c_void_p()
        _lib.check(_lib.lib().wgbv_open_slice(C.byref(props), first_node, node_count, halo_nodes, b.ctypes.data, b.size,
                                              o.ctypes.data, o.size, device, C.byref(h)))
        return cls(_Shared(h, None), cuda_stream)

    @classmethod
    def loadOnce(cls, stream):
        raise UnsupportedOperationException("GpuBVGraph needs random access to the bit stream")

    def close(self):
        if self._st is not None:
            self._st.close()
            self._st = None
            self._sh.release()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -------------------------------------------------------------------------------------------- ImmutableGraph
    def numNodes(self):
        return self._n

    def numArcs(self):
        return self._m

    def randomAccess(self):
        return True

    def hasCopiableIterators(self):
        return True

    def basename(self):
        return self._sh.basename

    def copy(self):
        """Flyweight copy (IG:447, BVG:553-578): shares the device-resident graph, owns its stream and scratch."""
        return GpuBVGraph(self._sh)

    def props(self):
        p = Props()
        _lib.check(_lib.lib().wgbv_get_props(self._sh.handle, C.byref(p)))
        return p

    def shardRange(self):
        return self._lo, self._hi

    def _check_node(self, x):
        if x < 0 or x >= self._n:
            raise IllegalArgumentException(f"Node index out of range: {x}")

    def outdegree(self, x):
        """BVGraph.outdegree(x), BVG:821-842."""
        self._check_node(x)
        return int(self.outdegrees(np.array([x], dtype=np.int64))[0])

    def outdegrees(self, ids):
        ids = np.ascontiguousarray(ids, dtype=np.int64)
        out = np.empty(ids.size, dtype=np.int64)
        _lib.check(_lib.lib().wgbv_outdegrees(self._st.h, ids.ctypes.data, ids.size, out.ctypes.data))
        return out

    def successorArray(self, x):
        """successorBigArray(x) (IG:329-333): the successors of x as an int64 array of length outdegree(x)."""
        self._check_node(x)
        _, col = self.successorsBatch(np.array([x], dtype=np.int64))
        return col

    successorBigArray = successorArray

    def successors(self, x):
        """BVGraph.successors(x), BVG:860-867: a lazy iterator ending with -1."""
        return LazyLongIterator(self.successorArray(x))

    def successorsBatch(self, ids):
        """Batched random access: (rowptr[k+1], col) for k arbitrary node ids (host arrays)."""
        ids = np.ascontiguousarray(ids, dtype=np.int64)
        L = _lib.lib()
        rowptr = np.empty(ids.size + 1, dtype=np.int64)
        arcs = C.c_int64()
        _lib.check(L.wgbv_successors_batch_host(self._st.h, ids.ctypes.data, ids.size, rowptr.ctypes.data, None, 0, C.byref(arcs)))
        col = np.empty(arcs.value, dtype=np.int64)
        _lib.check(L.wgbv_successors_batch_host(self._st.h, ids.ctypes.data, ids.size, rowptr.ctypes.data,
                                                col.ctypes.data if col.size else None, col.size, C.byref(arcs)))
        return rowptr, col

    def nodeIterator(self, from_=0):
        """BVGraph.nodeIterator(from), BVG:1257: sequential scan served from chunked bulk decodes."""
        if from_ < 0 or from_ > self._n:
            raise IllegalArgumentException(f"Node index out of range: {from_}")
        return NodeIterator(self, from_, self._n)

    def splitNodeIterators(self, howMany):
        """ImmutableGraph.splitNodeIterators(howMany), IG:405-436: contiguous ranges of ceil(n / howMany) nodes."""
        if self._n == 0 and howMany == 0:
            return []
        if howMany < 1:
            raise IllegalArgumentException()
        m = -(-self._n // howMany)
        res = []
        f = 0
        while f < self._n:
            res.append(NodeIterator(self.copy(), f, min(f + m, self._n), own_graph=True))
            f += m
        while len(res) < howMany:
            res.append(NodeIterator(self, self._n, self._n))
        return res

    # -------------------------------------------------------------------------------------------- bulk call
    def decodeToCSR(self, from_=None, to=None):
        """Bulk decode of [from, to) into host arrays: (rowptr int64[to-from+1], col int64[arcs])."""
        from_ = self._lo if from_ is None else from_
        to = self._hi if to is None else to
        L = _lib.lib()
        rowptr = np.empty(max(to - from_, 0) + 1, dtype=np.int64)
        arcs = C.c_int64()
        _lib.check(L.wgbv_decode_csr_host(self._st.h, from_, to, rowptr.ctypes.data, None, 0, C.byref(arcs)))
        col = np.empty(arcs.value, dtype=np.int64)
        _lib.check(L.wgbv_decode_csr_host(self._st.h, from_, to, rowptr.ctypes.data, col.ctypes.data if col.size else None,
                                          col.size, C.byref(arcs)))
        return rowptr, col

    def scan(self, from_=None, to=None, chunk_nodes=1 << 20, upload=None, narrow=False):
        """Streaming scan of [from, to): yields (first, last, rowptr, col) per chunk of nodes, rowptr relative to the chunk.
        The arrays are views of pinned buffers owned by the library and are only valid until the next chunk is requested
        (the reference's rule for iterator arrays, BVG:1171-1172); decode and copy-back of the following chunks overlap
        with the caller's work on the current one.  narrow=True (graphs of at most 2^32 nodes): col arrives as uint32,
        half the bytes over PCIe, same values."""
        from_ = self._lo if from_ is None else from_
        to = self._hi if to is None else to
        L = _lib.lib()
        _lib.check(L.wgbv_scan_set_narrow(self._st.h, 1 if narrow else 0))
        ctype, = (C.c_uint32,) if narrow else (C.c_int64,)
        if upload is None:
            _lib.check(L.wgbv_scan_begin(self._st.h, from_, to, chunk_nodes))
        else:  # (host pointer, bytes) of the compressed stream: uploaded piece by piece while the scan runs
            _lib.check(L.wgbv_scan_begin_upload(self._st.h, C.c_void_p(upload[0]), upload[1], from_, to, chunk_nodes))
        first, last, arcs = C.c_int64(), C.c_int64(), C.c_int64()
        rp, col = C.c_void_p(), C.c_void_p()
        try:
            while True:
                rc = L.wgbv_scan_next(self._st.h, C.byref(first), C.byref(last), C.byref(rp), C.byref(col), C.byref(arcs))
                if rc == 0:
                    return
                if rc < 0:
                    _lib.check(rc)
                n = last.value - first.value
                r = np.ctypeslib.as_array(C.cast(rp, C.POINTER(C.c_int64)), shape=(n + 1,))
                c = np.ctypeslib.as_array(C.cast(col, C.POINTER(ctype)), shape=(max(arcs.value, 1),))[:arcs.value]
                yield first.value, last.value, r, c
        finally:
            L.wgbv_scan_end(self._st.h)

    def decodeDegreesDevice(self, from_, to, d_rowptr):
        """Header pass + scan into a device buffer (int pointer); returns the arc count of [from, to)."""
        arcs = C.c_int64()
        _lib.check(_lib.lib().wgbv_decode_degrees(self._st.h, from_, to, C.c_void_p(d_rowptr), C.byref(arcs)))
        return arcs.value

    def decodeToCSRDevice(self, from_, to, d_rowptr, d_col, col_capacity, sync=True):
        """Bulk decode into caller-provided DEVICE buffers (raw pointers, e.g. torch tensors' data_ptr())."""
        _lib.check(_lib.lib().wgbv_decode_csr(self._st.h, from_, to, C.c_void_p(d_rowptr), C.c_void_p(d_col), col_capacity))
        if sync:
            self.sync()

    def sync(self):
        _lib.check(_lib.lib().wgbv_stream_sync(self._st.h))

    def uploadStream(self, host_ptr, nbytes):
        """Re-uploads the compressed bit stream from (pinned) host memory, asynchronously on this copy's stream."""
        _lib.check(_lib.lib().wgbv_upload_stream(self._st.h, C.c_void_p(host_ptr), nbytes))

    def decodeToCSRHostInto(self, from_, to, h_rowptr, h_col, col_capacity):
        """Bulk decode of [from, to) into caller-provided HOST buffers (raw pointers); returns the arc count."""
        arcs = C.c_int64()
        _lib.check(_lib.lib().wgbv_decode_csr_host(self._st.h, from_, to, C.c_void_p(h_rowptr), C.c_void_p(h_col),
                                                   col_capacity, C.byref(arcs)))
        return arcs.value

    def csrChecksumDevice(self, first_node, rows, d_rowptr, d_col):
        out = C.c_uint64()
        _lib.check(_lib.lib().wgbv_csr_checksum(self._st.h, first_node, rows, C.c_void_p(d_rowptr), C.c_void_p(d_col), C.byref(out)))
        return out.value

    def transposeCSRDevice(self, n, d_rowptr, d_col, d_t_rowptr, d_t_col):
        """Transpose of a CSR in device memory (int pointers; Transform.transposeOffline, Transform.java:1098-1144, on the
        device): d_t_rowptr gets n + 1 entries, d_t_col as many as d_col; sources of every target come out increasing."""
        _lib.check(_lib.lib().wgbv_transpose_csr(self._st.h, n, C.c_void_p(d_rowptr), C.c_void_p(d_col), C.c_void_p(d_t_rowptr),
                                                 C.c_void_p(d_t_col)))

    def symmetrizeCSRDevice(self, n, d_rowptr, d_col, d_s_rowptr, d_s_col=None, capacity=0):
        """Transform.symmetrizeOffline (union of the graph and its transpose, Transform.java:573-574) of a CSR in device memory;
        returns the arcs of the result (with d_s_col None only the row pointers and that count are produced)."""
        arcs = C.c_int64()
        _lib.check(_lib.lib().wgbv_symmetrize_csr(self._st.h, n, C.c_void_p(d_rowptr), C.c_void_p(d_col), C.c_void_p(d_s_rowptr),
                                                  C.c_void_p(d_s_col) if d_s_col else None, capacity, C.byref(arcs)))
        return arcs.value

    def unionCSRDevice(self, n, a_rowptr, a_col, b_rowptr, b_col, u_rowptr, u_col=None, capacity=0):
        """Transform.union of two CSRs over the same nodes, all in device memory; returns the arcs of the union."""
        arcs = C.c_int64()
        _lib.check(_lib.lib().wgbv_union_csr(self._st.h, n, C.c_void_p(a_rowptr), C.c_void_p(a_col), C.c_void_p(b_rowptr), C.c_void_p(b_col),
                                             C.c_void_p(u_rowptr), C.c_void_p(u_col) if u_col else None, capacity, C.byref(arcs)))
        return arcs.value

    def successorsBatchDevice(self, d_ids, k, d_rowptr, d_col=None, col_capacity=0):
        """Batched random access on device buffers.  With d_col None only the row pointers are produced."""
        L = _lib.lib()
        if d_col is None:
            arcs = C.c_int64()
            _lib.check(L.wgbv_successors_batch_degrees(self._st.h, C.c_void_p(d_ids), k, C.c_void_p(d_rowptr), C.byref(arcs)))
            return arcs.value
        _lib.check(L.wgbv_successors_batch(self._st.h, C.c_void_p(d_ids), k, C.c_void_p(d_rowptr), C.c_void_p(d_col), col_capacity))
        self.sync()
        return None

    def lastStageMs(self):
        out = (C.c_float * 5)()
        _lib.check(_lib.lib().wgbv_last_stage_ms(self._st.h, C.byref(out)))
        return list(out)

    def kernelLaunches(self):
        return _lib.lib().wgbv_kernel_launches(self._st.h)

    def fusedStats(self):
        """(tiles of the last bulk decode, tiles among them left to the slow-tile kernel, decodes rerun on the general
        path, whether this handle's decodes take the one-pass tile kernel); call after sync()."""
        out = (C.c_uint64 * 4)()
        _lib.check(_lib.lib().wgbv_fused_stats(self._st.h, C.byref(out)))
        return {"tiles": int(out[0]), "slow_tiles": int(out[1]), "reruns": int(out[2]), "enabled": bool(out[3])}

    def algorithmicBytes(self, from_, to, arcs):
        return _lib.lib().wgbv_algorithmic_bytes(self._sh.handle, from_, to, arcs)

    def deviceBytes(self):
        return _lib.lib().wgbv_device_bytes(self._sh.handle)

    def offsetsBytes(self):
        return _lib.lib().wgbv_offsets_bytes(self._sh.handle)

    # ImmutableGraph.equals, IG:759-779: same n; per node same outdegree and same successors in order
    def equals(self, other):
        if self.numNodes() != other.numNodes():
            return False
        a, b = self.nodeIterator(), other.nodeIterator()
        while a.hasNext():
            a.nextLong()
            b.nextLong()
            if a.outdegree() != b.outdegree() or not np.array_equal(a.successorArray(), b.successorArray()):
                return False
        return True


# coding ids of wgbv_props (CompressionFlags.java of the reference)
DELTA, GAMMA, GOLOMB, SKEWED_GOLOMB, UNARY, ZETA, NIBBLE = 1, 2, 3, 4, 5, 6, 7


def _gamma_stream(values):
    """The bytes of a gamma-coded stream of naturals (host side: the .offsets file of a stored graph)."""
    v = np.ascontiguousarray(values, dtype=np.uint64) + np.uint64(1)
    msb = np.zeros(v.size, dtype=np.int64)
    t = v.copy()
    for sh in (32, 16, 8, 4, 2, 1):
        big = t >= (np.uint64(1) << np.uint64(sh))
        msb[big] += sh
        t[big] >>= np.uint64(sh)
    ends = np.cumsum(2 * msb + 1)
    total = int(ends[-1]) if v.size else 0
    bits = np.zeros(total + 8, dtype=np.uint8)
    starts = ends - (2 * msb + 1)
    # code = msb zeros, then the msb + 1 significant bits of v
    for b in range(int(msb.max()) + 1 if v.size else 0):
        sel = msb >= b
        bits[(starts + msb + (msb - b))[sel]] = ((v[sel] >> np.uint64(b)) & np.uint64(1)).astype(np.uint8)
    return np.packbits(bits[: (total + 7) // 8 * 8]), total


def encode_csr_device(n, d_rowptr, d_col, arcs, window_size=7, max_ref_count=3, min_interval_length=4, zeta_k=3, residual_coding=ZETA,
                      first_node=0, split_nodes=0, device=0, out=None, want_node_bits=False):
    """wgbv_encode_csr on a CSR that is already in device memory (pointers): returns (graph bytes -- a view of `out` if
    given --, bits, record lengths or None)."""
    L = _lib.lib()
    p = _lib.Props(nodes=n, arcs=arcs, window_size=window_size, max_ref_count=max_ref_count, min_interval_length=min_interval_length, zeta_k=zeta_k,
                   outdegree_coding=GAMMA, block_coding=GAMMA, residual_coding=residual_coding, reference_coding=UNARY, block_count_coding=GAMMA,
                   offset_coding=GAMMA)
    if out is None:
        out = np.zeros(max(64, arcs * 10 + n * 4 + 64), dtype=np.uint8)
    node_bits = np.zeros(max(n, 1), dtype=np.uint64) if want_node_bits else None
    bits = C.c_uint64()
    _lib.check(L.wgbv_encode_csr(device, C.byref(p), first_node, n, C.c_void_p(d_rowptr), C.c_void_p(d_col), split_nodes, out.ctypes.data, out.size,
                                 C.byref(bits), node_bits.ctypes.data if want_node_bits else None))
    return out[: (bits.value + 7) // 8], bits.value, node_bits


def store_csr(rowptr, col, basename=None, window_size=7, max_ref_count=3, min_interval_length=4, zeta_k=3, residual_coding=ZETA,
              first_node=0, split_nodes=0, device=0):
    """BVGraph.store for successor lists given as a CSR (BVGraph.java:2329-2590): compresses on the device
    (wgbv_encode_csr) and returns (graph bytes, offsets bytes, properties text); with a basename the three files are
    written as well.  rowptr / col: numpy arrays, or device pointers given as (ptr, n) / (ptr,).  split_nodes > 0 starts
    every that many nodes with an empty window, like the threads of the reference's parallel compressor."""
    import torch
    L = _lib.lib()
    dev = torch.device("cuda", device)
    rp = torch.from_numpy(np.ascontiguousarray(rowptr, dtype=np.int64)).to(dev)
    cl = torch.from_numpy(np.ascontiguousarray(col, dtype=np.int64)).to(dev)
    n = rp.numel() - 1
    arcs = int(rowptr[-1] - rowptr[0])
    p = _lib.Props(nodes=n, arcs=arcs, window_size=window_size, max_ref_count=max_ref_count, min_interval_length=min_interval_length, zeta_k=zeta_k,
                   outdegree_coding=GAMMA, block_coding=GAMMA, residual_coding=residual_coding, reference_coding=UNARY, block_count_coding=GAMMA,
                   offset_coding=GAMMA)
    cap = max(64, arcs * 10 + n * 4 + 64)
    for _ in range(2):
        out = np.zeros(cap, dtype=np.uint8)
        node_bits = np.zeros(max(n, 1), dtype=np.uint64)
        bits = C.c_uint64()
        rc = L.wgbv_encode_csr(device, C.byref(p), first_node, n, C.c_void_p(rp.data_ptr()), C.c_void_p(cl.data_ptr() if cl.numel() else 0), split_nodes,
                               out.ctypes.data, out.size, C.byref(bits), node_bits.ctypes.data)
        if rc and bits.value // 8 + 1 > cap:
            cap = bits.value // 8 + 64
            continue
        _lib.check(rc)
        break
    graph = out[: (bits.value + 7) // 8].copy()
    offsets, _ = _gamma_stream(np.concatenate([np.zeros(1, dtype=np.uint64), node_bits[:n]]))
    flags = "" if residual_coding == ZETA else "RESIDUALS_" + {GAMMA: "GAMMA", DELTA: "DELTA", NIBBLE: "NIBBLE", GOLOMB: "GOLOMB"}[residual_coding]
    props = ("#BVGraph properties\ngraphclass=it.unimi.dsi.big.webgraph.BVGraph\nversion=0\n"
             f"nodes={n}\narcs={arcs}\nwindowsize={window_size}\nmaxrefcount={max_ref_count}\nminintervallength={min_interval_length}\n"
             f"zetak={zeta_k}\ncompressionflags={flags}\nbitsperlink={bits.value / max(arcs, 1):.4g}\nbitspernode={bits.value / max(n, 1):.4g}\n")
    if basename:
        with open(basename + ".graph", "wb") as f:
            f.write(graph.tobytes())
        with open(basename + ".offsets", "wb") as f:
            f.write(offsets.tobytes())
        with open(basename + ".properties", "w") as f:
            f.write(props)
    return graph, offsets, props


LABEL_FIXED_INT, LABEL_GAMMA_INT, LABEL_FIXED_INT_LIST, LABEL_FIXED_LONG_LIST = 0, 1, 2, 3


class GpuArcLabels:
    """The `.labels` / `.labeloffsets` pair of a BitStreamArcLabelledImmutableGraph on the device
    (labelling/BitStreamArcLabelledImmutableGraph.java:65-100): labels come back as an int64 array parallel to the
    successors of the underlying graph.  kind: LABEL_FIXED_INT (FixedWidthIntLabel, `width` bits), LABEL_GAMMA_INT, or the
    list-valued LABEL_FIXED_INT_LIST / LABEL_FIXED_LONG_LIST (decodeLists)."""

    def __init__(self, kind, width, labels_bytes, labeloffsets_bytes, nodes, device=0):
        lab = np.ascontiguousarray(np.frombuffer(labels_bytes, dtype=np.uint8)) if not isinstance(labels_bytes, np.ndarray) else np.ascontiguousarray(labels_bytes)
        off = np.ascontiguousarray(np.frombuffer(labeloffsets_bytes, dtype=np.uint8)) if not isinstance(labeloffsets_bytes, np.ndarray) else np.ascontiguousarray(labeloffsets_bytes)
        h = C.c_void_p()
        _lib.check(_lib.lib().wgbv_labels_open_mem(device, nodes, kind, width, lab.ctypes.data if lab.size else None, lab.size,
                                                   off.ctypes.data, off.size, C.byref(h)))
        self.h = h
        self.nodes = nodes

    def close(self):
        if self.h is not None:
            _lib.lib().wgbv_labels_close(self.h)
            self.h = None

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def decode(self, graph, rowptr, from_=0, to=None):
        """Labels of the arcs of nodes [from, to) (host arrays): rowptr[0 .. to-from] are the row pointers of that range,
        e.g. from graph.decodeToCSR(from, to); `graph` lends its stream."""
        to = self.nodes if to is None else to
        rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
        if rowptr.size != to - from_ + 1:
            raise IllegalArgumentException("rowptr must have to - from + 1 entries")
        arcs = int(rowptr[-1] - rowptr[0])
        out = np.empty(arcs, dtype=np.int64)
        _lib.check(_lib.lib().wgbv_labels_decode_host(graph._st.h, self.h, from_, to, rowptr.ctypes.data, out.ctypes.data if arcs else None, arcs))
        return out

    def decodeDevice(self, graph, from_, to, d_rowptr, d_labels, capacity):
        """Device buffers (int pointers), asynchronous on the graph copy's stream."""
        _lib.check(_lib.lib().wgbv_labels_decode(graph._st.h, self.h, from_, to, C.c_void_p(d_rowptr), C.c_void_p(d_labels), capacity))

    def decodeLists(self, graph, rowptr, from_=0, to=None):
        """List-valued labels (FixedWidthIntListLabel / FixedWidthLongListLabel) of the arcs of nodes [from, to): returns
        (list_ptr, values), arc j's list being values[list_ptr[j]:list_ptr[j + 1]] (host arrays)."""
        to = self.nodes if to is None else to
        rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
        if rowptr.size != to - from_ + 1:
            raise IllegalArgumentException("rowptr must have to - from + 1 entries")
        arcs = int(rowptr[-1] - rowptr[0])
        list_ptr = np.empty(arcs + 1, dtype=np.int64)
        total = C.c_int64(0)
        f = _lib.lib().wgbv_labels_decode_lists_host
        _lib.check(f(graph._st.h, self.h, from_, to, rowptr.ctypes.data, list_ptr.ctypes.data, None, 0, C.byref(total)))
        values = np.empty(max(total.value, 1), dtype=np.int64)
        _lib.check(f(graph._st.h, self.h, from_, to, rowptr.ctypes.data, list_ptr.ctypes.data, values.ctypes.data, total.value, C.byref(total)))
        return list_ptr, values[: total.value]

    def decodeListsDevice(self, graph, from_, to, d_rowptr, d_list_ptr, d_values, capacity):
        """Device buffers; d_values 0: only the list starts.  Returns the number of elements of the range."""
        total = C.c_int64(0)
        _lib.check(_lib.lib().wgbv_labels_decode_lists(graph._st.h, self.h, from_, to, C.c_void_p(d_rowptr), C.c_void_p(d_list_ptr),
                                                       C.c_void_p(d_values) if d_values else None, capacity, C.byref(total)))
        return total.value


class NodeIterator:
    """BVGraphNodeIterator (BVG:1100-1245) over chunked GPU decodes; arrays returned are valid until the next
    nextLong() (the reference's window-slot rule, BVG:1171-1172)."""

    def __init__(self, graph, from_, upper_bound, own_graph=False):
        self._g = graph
        self._own = own_graph
        self._from = from_
        self._curr = from_ - 1
        self._limit = min(upper_bound, graph.numNodes()) - 1  # hasNextLimit, BVG:1147
        self._c0 = self._c1 = from_  # chunk [c0, c1) currently held
        self._rowptr = self._col = None

    def hasNext(self):
        return self._curr < self._limit

    def nextLong(self):
        if not self.hasNext():
            raise NoSuchElementException()
        self._curr += 1
        if self._curr >= self._c1:
            self._c0 = self._curr
            self._c1 = min(self._c0 + GpuBVGraph.CHUNK_NODES, self._limit + 1)
            self._rowptr, self._col = self._g.decodeToCSR(self._c0, self._c1)
        return self._curr

    __next__ = nextLong

    def __iter__(self):
        while self.hasNext():
            yield self.nextLong()

    def _row(self):
        if self._curr == self._from - 1:
            raise IllegalStateException()
        i = self._curr - self._c0
        return self._rowptr[i], self._rowptr[i + 1]

    def outdegree(self):
        a, b = self._row()
        return int(b - a)

    def successorArray(self):
        a, b = self._row()
        return self._col[a:b]

    successorBigArray = successorArray

    def successors(self):
        return LazyLongIterator(self.successorArray())

    def skip(self, n):
        if n < 0:
            raise IllegalArgumentException(f"Argument must be nonnegative: {n}")
        k = 0
        while k < n and self.hasNext():
            self.nextLong()
            k += 1
        return k

    def copy(self, upper_bound):
        """NodeIterator.copy(upperBound), BVG:1223-1229: an independent iterator over [curr+1, upperBound)."""
        return NodeIterator(self._g.copy(), self._curr + 1, upper_bound, own_graph=True)
