"""ctypes front end of datagen/bvgen.c: synthetic graphs BV-encoded the way the reference's compressor would.

Produces in-memory ``.graph`` / ``.offsets`` / ``.properties`` triples (and can write them to disk) plus
an independent checksum of the *uncompressed* successor lists taken before encoding, which the CUDA
decoder's checksum kernel has to reproduce.  Inputs only: this is not part of the decode path.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

DELTA, GAMMA, GOLOMB, UNARY, ZETA, NIBBLE = 1, 2, 3, 5, 6, 7
MODEL_CSR, MODEL_ER, MODEL_COPY, MODEL_BIG = 0, 1, 2, 3
_NAMES = {1: "DELTA", 2: "GAMMA", 3: "GOLOMB", 5: "UNARY", 6: "ZETA", 7: "NIBBLE"}


class Params(C.Structure):
    _fields_ = [("window", C.c_int32), ("maxref", C.c_int32), ("minint", C.c_int32), ("zetak", C.c_int32),
                ("outdegree_coding", C.c_int32), ("block_coding", C.c_int32), ("residual_coding", C.c_int32),
                ("reference_coding", C.c_int32), ("blockcount_coding", C.c_int32), ("offset_coding", C.c_int32)]


class Model(C.Structure):
    _fields_ = [("model", C.c_int32), ("n", C.c_int64), ("first_node", C.c_int64), ("global_n", C.c_int64),
                ("seed", C.c_uint64), ("p", C.c_double),
                ("mean_degree", C.c_double), ("p_ref", C.c_double), ("p_keep", C.c_double), ("p_run", C.c_double),
                ("far_frac", C.c_double), ("near_scale", C.c_double), ("max_degree", C.c_int64),
                ("giant_every", C.c_int64), ("giant_degree", C.c_int64), ("giant_stride", C.c_int64),
                ("rowptr", C.c_void_p), ("col", C.c_void_p)]


def build(force=False):
    so = os.path.join(_HERE, "libbvgen.so")
    src = os.path.join(_HERE, "bvgen.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "libbvgen.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        L.bvg_run.argtypes = [C.POINTER(Params), C.POINTER(Model), C.c_int, C.c_int, C.c_int64]
        L.bvg_run.restype = C.c_void_p
        L.bvg_graph.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.bvg_graph.restype = C.c_void_p
        L.bvg_offsets.argtypes = [C.c_void_p]
        L.bvg_offsets.restype = C.c_void_p
        L.bvg_offsets_stream.argtypes = [C.c_void_p, C.POINTER(C.c_uint64)]
        L.bvg_offsets_stream.restype = C.c_void_p
        L.bvg_arcs.argtypes = [C.c_void_p]
        L.bvg_arcs.restype = C.c_int64
        L.bvg_stats.argtypes = [C.c_void_p, C.c_void_p]
        L.bvg_stats.restype = None
        L.bvg_error.argtypes = [C.c_void_p]
        L.bvg_error.restype = C.c_int
        L.bvg_checksum.argtypes = [C.c_void_p]
        L.bvg_checksum.restype = C.c_uint64
        L.bvg_seg_checksums.argtypes = [C.c_void_p, C.POINTER(C.c_int64)]
        L.bvg_seg_checksums.restype = C.c_void_p
        L.bvg_free.argtypes = [C.c_void_p]
        L.bvg_free.restype = None
        L.bvg_csr_checksum.argtypes = [C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
        L.bvg_csr_checksum.restype = C.c_uint64
        L.bvg_write_labels.argtypes = [C.c_int, C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
        L.bvg_write_labels.restype = C.c_uint64
        L.bvg_write_label_lists.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
        L.bvg_write_label_lists.restype = C.c_uint64
        L.bvg_write_labels_synthetic.argtypes = [C.c_int, C.c_int64, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]
        L.bvg_write_labels_synthetic.restype = C.c_uint64
        L.bvg_write_codes.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_int64, C.c_void_p, C.c_uint64]
        L.bvg_write_codes.restype = C.c_uint64
        _LIB = L
    return _LIB


def make_params(window=7, maxref=3, minint=4, zetak=3, outdegree=GAMMA, block=GAMMA, residual=ZETA,
                reference=UNARY, blockcount=GAMMA, offset=GAMMA):
    return Params(window, maxref, minint, zetak, outdegree, block, residual, reference, blockcount, offset)


def flags_string(p):
    """compressionflags value for a parameter set: only non-default codings are named (BVG:1281-1315)."""
    out = []
    for field, default, prefix in (("outdegree_coding", GAMMA, "OUTDEGREES"), ("block_coding", GAMMA, "BLOCKS"),
                                   ("residual_coding", ZETA, "RESIDUALS"), ("reference_coding", UNARY, "REFERENCES"),
                                   ("blockcount_coding", GAMMA, "BLOCK_COUNT"), ("offset_coding", GAMMA, "OFFSETS")):
        v = getattr(p, field)
        if v != default:
            out.append(f"{prefix}_{_NAMES[v]}")
    return " | ".join(out)


# what the reference's compressor writes to .properties (BVGraph.java:2490-2567), plus a few shape figures of our own
STAT_NAMES = ("copiedarcs", "intervalisedarcs", "residualarcs", "rows_with_reference", "blocks", "intervals", "bitsforoutdegrees",
              "bitsforreferences", "bitsforblocks", "bitsforintervals", "bitsforresiduals", "rows_with_reference_row_over_128",
              "refchain_length_sum", "rows_chain1", "rows_chain2", "rows_chain3", "rows_chain4plus")


class Encoded:
    """An encoded graph held in host memory (numpy arrays own the bytes)."""

    def __init__(self, params, n, arcs, graph, graph_bits, offsets, offsets_stream, checksum, seg_checksums, seg_nodes, stats=None):
        self.params, self.n, self.arcs = params, n, arcs
        self.stats = stats or {}
        self.graph, self.graph_bits = graph, graph_bits
        self.offsets, self.offsets_stream = offsets, offsets_stream
        self.checksum, self.seg_checksums, self.seg_nodes = checksum, seg_checksums, seg_nodes

    def properties_text(self):
        p = self.params
        return ("#BVGraph properties\n"
                "graphclass=it.unimi.dsi.big.webgraph.BVGraph\n"
                "version=0\n"
                f"nodes={self.n}\narcs={self.arcs}\n"
                f"windowsize={p.window}\nmaxrefcount={p.maxref}\nminintervallength={p.minint}\nzetak={p.zetak}\n"
                f"compressionflags={flags_string(p)}\n")

    def write(self, basename):
        with open(basename + ".graph", "wb") as f:
            f.write(self.graph.tobytes())
        with open(basename + ".offsets", "wb") as f:
            f.write(self.offsets_stream.tobytes())
        with open(basename + ".properties", "w") as f:
            f.write(self.properties_text())
        return basename


def _collect(job, params, n, seg_nodes, want_offsets_stream=True):
    L = lib()
    try:
        if L.bvg_error(job):
            raise RuntimeError("bvgen: a node record exceeded 2^32-1 bits")
        nbytes, nbits = C.c_uint64(), C.c_uint64()
        gp = L.bvg_graph(job, C.byref(nbytes), C.byref(nbits))
        graph = np.ctypeslib.as_array(C.cast(gp, C.POINTER(C.c_uint8)), shape=(max(nbytes.value, 1),))[:nbytes.value].copy()
        offsets = np.ctypeslib.as_array(C.cast(L.bvg_offsets(job), C.POINTER(C.c_uint64)), shape=(n + 1,)).copy()
        ostream = None
        if want_offsets_stream:
            ob = C.c_uint64()
            op = L.bvg_offsets_stream(job, C.byref(ob))
            ostream = np.ctypeslib.as_array(C.cast(op, C.POINTER(C.c_uint8)), shape=(ob.value,)).copy()
        segs = None
        if seg_nodes:
            ns = C.c_int64()
            sp = L.bvg_seg_checksums(job, C.byref(ns))
            segs = np.ctypeslib.as_array(C.cast(sp, C.POINTER(C.c_uint64)), shape=(max(ns.value, 1),))[:ns.value].copy()
        st = np.zeros(20, dtype=np.int64)
        L.bvg_stats(job, st.ctypes.data)
        stats = {k: int(v) for k, v in zip(STAT_NAMES, st)}
        return Encoded(params, n, L.bvg_arcs(job), graph, nbits.value, offsets, ostream, L.bvg_checksum(job), segs, seg_nodes, stats)
    finally:
        L.bvg_free(job)


def _threads(threads):
    return threads if threads else max(1, len(os.sched_getaffinity(0)))


def encode_csr(rowptr, col, params=None, nsplits=1, threads=0, seg_nodes=0):
    params = params or make_params()
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
    col = np.ascontiguousarray(col, dtype=np.int64)
    n = rowptr.size - 1
    m = Model(model=MODEL_CSR, n=n, rowptr=rowptr.ctypes.data, col=col.ctypes.data if col.size else None)
    job = lib().bvg_run(C.byref(params), C.byref(m), nsplits, _threads(threads), seg_nodes)
    return _collect(job, params, n, seg_nodes)


def erdos_renyi(n, p, seed=1, params=None, nsplits=64, threads=0, seg_nodes=0):
    params = params or make_params()
    m = Model(model=MODEL_ER, n=n, seed=seed, p=p)
    return _collect(lib().bvg_run(C.byref(params), C.byref(m), nsplits, _threads(threads), seg_nodes), params, n, seg_nodes)


def copy_model(n, seed=2, mean_degree=20.0, p_ref=0.7, p_keep=0.8, p_run=0.3, far_frac=0.05, near_scale=1e4,
               max_degree=10000, params=None, nsplits=256, threads=0, seg_nodes=0, first_node=0, global_n=0,
               giant_every=0, giant_degree=1 << 20, giant_stride=4):
    """Copy-model web-like graph; with first_node/global_n one part [first_node, first_node+n) of a larger graph.
    giant_every > 0: node x with x % giant_every == giant_every // 2 gets giant_degree successors i * giant_stride instead
    (the giant rows of the reference's slow test, BVGraphSlowTest.java:30-96)."""
    params = params or make_params()
    m = Model(model=MODEL_COPY, n=n, first_node=first_node, global_n=global_n, seed=seed, mean_degree=mean_degree, p_ref=p_ref, p_keep=p_keep, p_run=p_run,
              far_frac=far_frac, near_scale=near_scale, max_degree=max_degree, giant_every=giant_every, giant_degree=giant_degree,
              giant_stride=giant_stride)
    return _collect(lib().bvg_run(C.byref(params), C.byref(m), nsplits, _threads(threads), seg_nodes), params, n, seg_nodes)


def big_graph(n, giant_every=0, giant_degree=1 << 20, giant_stride=4, params=None, nsplits=64, threads=0, seg_nodes=0):
    params = params or make_params()
    m = Model(model=MODEL_BIG, n=n, giant_every=giant_every, giant_degree=giant_degree, giant_stride=giant_stride)
    return _collect(lib().bvg_run(C.byref(params), C.byref(m), nsplits, _threads(threads), seg_nodes), params, n, seg_nodes)


def csr_checksum(first_node, rowptr, col):
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
    col = np.ascontiguousarray(col, dtype=np.int64)
    return lib().bvg_csr_checksum(first_node, rowptr.size - 1, rowptr.ctypes.data, col.ctypes.data if col.size else None)


def write_codes(coding, values, k=3):
    vals = np.ascontiguousarray(values, dtype=np.uint64)
    per = 17 if coding not in (UNARY, GOLOMB) else int(vals.max() if vals.size else 0) // 8 + 17
    out = np.zeros(vals.size * per + 64, dtype=np.uint8)
    bits = lib().bvg_write_codes(coding, k, vals.ctypes.data, vals.size, out.ctypes.data, out.size)
    return out[: (bits + 7) // 8].copy(), bits


LABEL_FIXED_INT, LABEL_GAMMA_INT, LABEL_FIXED_INT_LIST, LABEL_FIXED_LONG_LIST = 0, 1, 2, 3


def write_labels(kind, width, rowptr, values):
    """The .labels and .labeloffsets streams of a labelled graph whose arcs (in CSR order) carry `values`
    (BitStreamArcLabelledImmutableGraph.store restated, see bvg_write_labels).  Returns (labels bytes, labeloffsets bytes)."""
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
    values = np.ascontiguousarray(values, dtype=np.int64)
    n = rowptr.size - 1
    per = width if kind == LABEL_FIXED_INT else 2 * max(int(values.max()) if values.size else 0, 1).bit_length() + 1
    out = np.zeros(values.size * per // 8 + 64, dtype=np.uint8)
    node_bits = np.zeros(max(n, 1), dtype=np.uint64)
    bits = lib().bvg_write_labels(kind, width, n, rowptr.ctypes.data, values.ctypes.data if values.size else None, out.ctypes.data, out.size,
                                  node_bits.ctypes.data)
    assert (bits + 7) // 8 <= out.size
    offs, _ = write_codes(GAMMA, np.concatenate([np.zeros(1, dtype=np.uint64), node_bits[:n]]))
    return out[: (bits + 7) // 8].copy(), offs


def write_label_lists(width, rowptr, list_ptr, values):
    """The .labels and .labeloffsets streams of a graph whose arc j (CSR order) carries the list
    values[list_ptr[j]:list_ptr[j + 1]] of `width`-bit elements (see bvg_write_label_lists)."""
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
    list_ptr = np.ascontiguousarray(list_ptr, dtype=np.int64)
    values = np.ascontiguousarray(values, dtype=np.int64)
    n = rowptr.size - 1
    longest = int(np.diff(list_ptr).max()) if list_ptr.size > 1 else 0
    out = np.zeros((values.size * width + (list_ptr.size - 1) * (2 * (longest + 1).bit_length() + 1)) // 8 + 64, dtype=np.uint8)
    node_bits = np.zeros(max(n, 1), dtype=np.uint64)
    bits = lib().bvg_write_label_lists(width, n, rowptr.ctypes.data, list_ptr.ctypes.data, values.ctypes.data if values.size else None,
                                       out.ctypes.data, out.size, node_bits.ctypes.data)
    assert (bits + 7) // 8 <= out.size
    offs, _ = write_codes(GAMMA, np.concatenate([np.zeros(1, dtype=np.uint64), node_bits[:n]]))
    return out[: (bits + 7) // 8].copy(), offs


LABEL_HASH_MULTIPLIER = 0x9E3779B97F4A7C15


def write_labels_synthetic(width, rowptr):
    """Fixed-width labels for a benchmark-size graph without a value array: arc j (CSR order) carries the low `width`
    bits of (j * LABEL_HASH_MULTIPLIER mod 2^64) >> 20.  Returns (labels bytes, labeloffsets bytes)."""
    rowptr = np.ascontiguousarray(rowptr, dtype=np.int64)
    n = rowptr.size - 1
    arcs = int(rowptr[-1] - rowptr[0])
    out = np.zeros((arcs * width + 7) // 8 + 64, dtype=np.uint8)
    node_bits = np.zeros(max(n, 1), dtype=np.uint64)
    bits = lib().bvg_write_labels_synthetic(width, n, rowptr.ctypes.data, out.ctypes.data, out.size, node_bits.ctypes.data)
    assert bits == arcs * width
    offs, _ = write_codes(GAMMA, np.concatenate([np.zeros(1, dtype=np.uint64), node_bits[:n]]))
    return out[: (bits + 7) // 8], offs
