"""ctypes front end of the CPU oracle (oracle/bvoracle.c).

TEST INFRASTRUCTURE ONLY: imported by tests/, by __graft_entry__.smoke() and by bench.py's cpu_baseline /
``--impl reference`` legs -- never by the product package.  See the header of bvoracle.c for the parity status.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

# CompressionFlags.java:26-46
DELTA, GAMMA, GOLOMB, SKEWED_GOLOMB, UNARY, ZETA, NIBBLE = 1, 2, 3, 4, 5, 6, 7
_CODING = {"DELTA": DELTA, "GAMMA": GAMMA, "GOLOMB": GOLOMB, "SKEWED_GOLOMB": SKEWED_GOLOMB,
           "UNARY": UNARY, "ZETA": ZETA, "NIBBLE": NIBBLE}
# bit slot of each field inside the flag mask, BVG:475-524 / setFlags BVG:1281-1289
_SLOT = {"OUTDEGREES": 0, "BLOCKS": 4, "RESIDUALS": 8, "REFERENCES": 12, "BLOCK_COUNT": 16, "OFFSETS": 20}
# the constants BVGraph actually declares (string2Flags resolves names by reflection, BVG:1316-1331)
_DECLARED = {
    "OUTDEGREES_GAMMA", "OUTDEGREES_DELTA", "BLOCKS_GAMMA", "BLOCKS_DELTA", "RESIDUALS_GAMMA", "RESIDUALS_ZETA",
    "RESIDUALS_DELTA", "RESIDUALS_NIBBLE", "RESIDUALS_GOLOMB", "REFERENCES_GAMMA", "REFERENCES_DELTA",
    "REFERENCES_UNARY", "BLOCK_COUNT_GAMMA", "BLOCK_COUNT_DELTA", "BLOCK_COUNT_UNARY", "OFFSETS_GAMMA",
    "OFFSETS_DELTA",
}


class Params(C.Structure):
    _fields_ = [("nodes", C.c_int64), ("arcs", C.c_int64),
                ("window", C.c_int32), ("maxref", C.c_int32), ("minint", C.c_int32), ("zetak", C.c_int32),
                ("outdegree_coding", C.c_int32), ("block_coding", C.c_int32), ("residual_coding", C.c_int32),
                ("reference_coding", C.c_int32), ("blockcount_coding", C.c_int32), ("offset_coding", C.c_int32)]


def build(force=False):
    so = os.path.join(_HERE, "libbvoracle.so")
    src = os.path.join(_HERE, "bvoracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s", "-B", "libbvoracle.so"])
    return so


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(build())
        u8p, u64p, i64p = C.POINTER(C.c_uint8), C.POINTER(C.c_uint64), C.POINTER(C.c_int64)
        L.bvo_read_offsets.argtypes = [C.POINTER(Params), C.c_void_p, C.c_uint64, C.c_void_p]
        L.bvo_read_offsets.restype = C.c_int
        L.bvo_outdegree.argtypes = [C.POINTER(Params), C.c_void_p, C.c_uint64, C.c_void_p, C.c_int64]
        L.bvo_outdegree.restype = C.c_int64
        L.bvo_successors.argtypes = [C.POINTER(Params), C.c_void_p, C.c_uint64, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64]
        L.bvo_successors.restype = C.c_int64
        L.bvo_decode_range.argtypes = [C.POINTER(Params), C.c_void_p, C.c_uint64, C.c_void_p, C.c_int64, C.c_int64,
                                       C.c_void_p, C.c_void_p, C.c_int64]
        L.bvo_decode_range.restype = C.c_int64
        L.bvo_decode_split.argtypes = [C.POINTER(Params), C.c_void_p, C.c_uint64, C.c_void_p, C.c_int64, C.c_int64,
                                       C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        L.bvo_decode_split.restype = C.c_int64
        L.bvo_decode_labels.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
        L.bvo_decode_labels.restype = C.c_int64
        L.bvo_decode_label_lists.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_uint64, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_int64]
        L.bvo_decode_label_lists.restype = C.c_int64
        L.bvo_random_access.argtypes = [C.POINTER(Params), C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_int64, C.c_int]
        L.bvo_random_access.restype = C.c_int64
        L.bvo_read_code.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64), C.c_int, C.c_int]
        L.bvo_read_code.restype = C.c_uint64
        del u8p, u64p, i64p
        _LIB = L
    return _LIB


def parse_properties(text):
    """java.util.Properties subset: key=value / key:value lines, # and ! comments."""
    props = {}
    for line in text.splitlines():
        line = line.strip()
        if not line or line[0] in "#!":
            continue
        for i, ch in enumerate(line):
            if ch in "=:":
                props[line[:i].strip()] = line[i + 1:].strip()
                break
        else:
            props[line] = ""
    return props


def string2flags(s):
    """BVGraph.string2Flags, BVG:1316-1331."""
    flags = 0
    if s:
        for name in s.split("|"):
            name = name.strip()
            if name not in _DECLARED:
                raise IOError(f"Compression flag {name} unknown.")
            field, coding = next((f, name[len(f) + 1:]) for f in _SLOT if name.startswith(f + "_"))
            flags |= _CODING[coding] << _SLOT[field]
    return flags


def params_from_properties(props):
    """The checks and assignments of BVGraph.loadInternal, BVG:1482-1503."""
    gc = props.get("graphclass", "")
    if gc.replace("it.unimi.dsi.webgraph", "it.unimi.dsi.big.webgraph") != "it.unimi.dsi.big.webgraph.BVGraph":
        raise IOError(f'cannot load a graph stored using class "{gc}"')
    if "version" not in props:
        raise IOError("Missing format version information")
    if int(props["version"]) > 0:
        raise IOError("format too new")
    flags = string2flags(props.get("compressionflags", ""))
    p = Params(nodes=int(props["nodes"]), arcs=int(props["arcs"]), window=int(props["windowsize"]),
               maxref=int(props["maxrefcount"]), minint=int(props["minintervallength"]),
               zetak=int(props.get("zetak", 3)),
               outdegree_coding=GAMMA, block_coding=GAMMA, residual_coding=ZETA, reference_coding=UNARY,
               blockcount_coding=GAMMA, offset_coding=GAMMA)
    for slot, field in ((0, "outdegree_coding"), (4, "block_coding"), (8, "residual_coding"),
                        (12, "reference_coding"), (16, "blockcount_coding"), (20, "offset_coding")):
        v = (flags >> slot) & 0xF
        if v:
            setattr(p, field, v)
    return p


class OracleGraph:
    """A BVGraph held in host memory and decoded by the CPU oracle."""

    def __init__(self, params, graph_bytes, offsets_bytes=None, offsets=None):
        self.p = params
        self.graph = np.frombuffer(graph_bytes, dtype=np.uint8) if not isinstance(graph_bytes, np.ndarray) else graph_bytes
        self.graph = np.ascontiguousarray(self.graph)
        if offsets is None and offsets_bytes is not None:
            ob = np.ascontiguousarray(np.frombuffer(offsets_bytes, dtype=np.uint8))
            offsets = np.zeros(params.nodes + 1, dtype=np.uint64)
            rc = lib().bvo_read_offsets(C.byref(self.p), ob.ctypes.data, ob.size, offsets.ctypes.data)
            if rc:
                raise IOError(f"bvo_read_offsets failed: {rc}")
        self.offsets = offsets

    @classmethod
    def load(cls, basename):
        with open(basename + ".properties") as f:
            params = params_from_properties(parse_properties(f.read()))
        with open(basename + ".graph", "rb") as f:
            g = f.read()
        with open(basename + ".offsets", "rb") as f:
            o = f.read()
        return cls(params, g, o)

    @property
    def num_nodes(self):
        return self.p.nodes

    def _off(self):
        return self.offsets.ctypes.data if self.offsets is not None else None

    def outdegree(self, x):
        d = lib().bvo_outdegree(C.byref(self.p), self.graph.ctypes.data, self.graph.size, self._off(), x)
        if d < 0:
            raise _err(d)
        return d

    def successors(self, x):
        d = self.outdegree(x)
        out = np.empty(max(d, 1), dtype=np.int64)
        r = lib().bvo_successors(C.byref(self.p), self.graph.ctypes.data, self.graph.size, self._off(), x, out.ctypes.data, out.size)
        if r < 0:
            raise _err(r)
        return out[:r]

    def random_access(self, ids, threads=1):
        """successors(x) for every x of ids on `threads` host threads (SpeedTest -r); returns the arcs read."""
        ids = np.ascontiguousarray(ids, dtype=np.int64)
        r = lib().bvo_random_access(C.byref(self.p), self.graph.ctypes.data, self.graph.size, self._off(), ids.ctypes.data, ids.size, threads)
        if r < 0:
            raise _err(r)
        return r

    def decode_range(self, lo=0, hi=None, threads=1, want_col=True):
        """Sequential (threads == 1, nodeIterator) or split (splitNodeIterators) scan of [lo, hi): (deg, col)."""
        hi = self.p.nodes if hi is None else hi
        n = max(hi - lo, 0)
        deg = np.zeros(n, dtype=np.int64)
        L = lib()
        g, gb = self.graph.ctypes.data, self.graph.size
        if threads <= 1:
            arcs = L.bvo_decode_range(C.byref(self.p), g, gb, self._off(), lo, hi, deg.ctypes.data, None, 0)
            if arcs < 0:
                raise _err(arcs)
            if not want_col:
                return deg, None
            col = np.empty(max(arcs, 1), dtype=np.int64)
            arcs2 = L.bvo_decode_range(C.byref(self.p), g, gb, self._off(), lo, hi, deg.ctypes.data, col.ctypes.data, col.size)
            assert arcs2 == arcs, (arcs, arcs2)
            return deg, col[:arcs]
        arcs = L.bvo_decode_split(C.byref(self.p), g, gb, self._off(), lo, hi, threads, deg.ctypes.data, None, None)
        if arcs < 0:
            raise _err(arcs)
        if not want_col:
            return deg, None
        rowptr = np.zeros(n + 1, dtype=np.int64)
        np.cumsum(deg, out=rowptr[1:])
        col = np.empty(max(arcs, 1), dtype=np.int64)
        arcs2 = L.bvo_decode_split(C.byref(self.p), g, gb, self._off(), lo, hi, threads, deg.ctypes.data, col.ctypes.data, rowptr.ctypes.data)
        assert arcs2 == arcs
        return deg, col[:arcs]


def read_code(buf, bitpos, coding, k=3):
    b = np.ascontiguousarray(np.frombuffer(buf, dtype=np.uint8))
    pos = C.c_uint64(bitpos)
    v = lib().bvo_read_code(b.ctypes.data, b.size, C.byref(pos), coding, k)
    return v, pos.value


def _err(code):
    return {-1: IndexError("Node index out of range"), -2: RuntimeError("illegal state (malformed stream / bad reference)"),
            -3: IOError("bit stream overrun"), -4: NotImplementedError("unsupported coding"),
            -5: MemoryError("capacity")}.get(code, RuntimeError(f"oracle error {code}"))


def transpose_csr(n, rowptr, col):
    """CPU restatement of what Transform.transposeOffline / Transform.transpose compute (Transform.java:1098-1144:
    every arc (x, y) of the scan becomes (y, x), arcs are sorted by (target, source) in batches and merged): returns
    (t_rowptr, t_col) with the sources of every target increasing.  TEST INFRASTRUCTURE ONLY, like the rest of oracle/.
    Parity status: unpinned by the reference (its tests only check transpose(transpose(g)) == g, TransformTest); pinned
    here by that involution on the cnr-2000 golden and by this definition."""
    rowptr = np.asarray(rowptr, dtype=np.int64)
    col = np.asarray(col, dtype=np.int64)
    src = np.repeat(np.arange(n, dtype=np.int64), np.diff(rowptr))
    order = np.lexsort((src, col))  # by target, then by source
    t_rowptr = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(col, minlength=n)[:n], out=t_rowptr[1:])
    return t_rowptr, src[order]


def union_csr(n, a_rowptr, a_col, b_rowptr, b_col):
    """CPU restatement of Transform.union (Transform.java:1659-1663, UnionImmutableGraph): per node the increasing union of
    the two successor lists, duplicates dropped.  TEST INFRASTRUCTURE ONLY.  Parity status: unpinned by the reference (no
    stored fixture of a union); pinned here by this definition."""
    src = np.concatenate([np.repeat(np.arange(n, dtype=np.int64), np.diff(a_rowptr)), np.repeat(np.arange(n, dtype=np.int64), np.diff(b_rowptr))])
    dst = np.concatenate([np.asarray(a_col, dtype=np.int64), np.asarray(b_col, dtype=np.int64)])
    pairs = np.unique(np.stack([src, dst], axis=1), axis=0) if src.size else np.zeros((0, 2), dtype=np.int64)
    rp = np.zeros(n + 1, dtype=np.int64)
    np.cumsum(np.bincount(pairs[:, 0], minlength=n)[:n], out=rp[1:])
    return rp, pairs[:, 1].astype(np.int64)


def symmetrize_csr(n, rowptr, col):
    """Transform.symmetrizeOffline restated (Transform.java:573-574): union(g, transpose(g))."""
    trp, tcol = transpose_csr(n, rowptr, col)
    return union_csr(n, rowptr, col, trp, tcol)


LABEL_FIXED_INT, LABEL_GAMMA_INT, LABEL_FIXED_INT_LIST, LABEL_FIXED_LONG_LIST = 0, 1, 2, 3


def _label_offsets(labeloffsets_bytes, n):
    offs = np.ascontiguousarray(np.frombuffer(labeloffsets_bytes, dtype=np.uint8))
    p = Params()
    p.nodes = n
    p.offset_coding = 2  # BVO_GAMMA
    off = np.zeros(n + 1, dtype=np.uint64)
    rc = lib().bvo_read_offsets(C.byref(p), offs.ctypes.data, offs.size, off.ctypes.data)
    if rc:
        raise _err(rc)
    return off


def decode_label_lists(kind, width, labels_bytes, labeloffsets_bytes, degrees, lo=0, hi=None):
    """List-valued labels of the arcs of nodes [lo, hi) (see bvo_decode_label_lists): returns (list_ptr, values) with
    arc j's list = values[list_ptr[j]:list_ptr[j + 1]].  Parity status: unpinned by the reference, like decode_labels."""
    degrees = np.ascontiguousarray(degrees, dtype=np.int64)
    n = degrees.size
    hi = n if hi is None else hi
    lab = np.ascontiguousarray(np.frombuffer(labels_bytes, dtype=np.uint8))
    off = _label_offsets(labeloffsets_bytes, n)
    part = np.ascontiguousarray(degrees[lo:hi])
    lens = np.zeros(int(part.sum()), dtype=np.int64)
    is_long = int(kind == LABEL_FIXED_LONG_LIST)
    total = lib().bvo_decode_label_lists(is_long, width, lab.ctypes.data, lab.size, off.ctypes.data, lo, hi, part.ctypes.data, lens.ctypes.data, None, 0)
    if total < 0:
        raise ValueError("label stream and label offsets disagree")
    values = np.zeros(total, dtype=np.int64)
    got = lib().bvo_decode_label_lists(is_long, width, lab.ctypes.data, lab.size, off.ctypes.data, lo, hi, part.ctypes.data, lens.ctypes.data,
                                       values.ctypes.data, total)
    assert got == total
    list_ptr = np.zeros(lens.size + 1, dtype=np.int64)
    np.cumsum(lens, out=list_ptr[1:])
    return list_ptr, values


def decode_labels(kind, width, labels_bytes, labeloffsets_bytes, degrees, lo=0, hi=None):
    """Labels of the arcs of nodes [lo, hi), in arc order (BitStreamArcLabelledImmutableGraph restated, see
    bvo_decode_labels): degrees[i] is the outdegree of node i of the WHOLE graph.  Parity status: unpinned by the
    reference (its labelled-graph tests generate their inputs with Java code at run time; there is no stored fixture)."""
    degrees = np.ascontiguousarray(degrees, dtype=np.int64)
    n = degrees.size
    hi = n if hi is None else hi
    lab = np.ascontiguousarray(np.frombuffer(labels_bytes, dtype=np.uint8))
    offs = np.ascontiguousarray(np.frombuffer(labeloffsets_bytes, dtype=np.uint8))
    p = Params()
    p.nodes = n
    p.offset_coding = 2  # BVO_GAMMA
    off = np.zeros(n + 1, dtype=np.uint64)
    rc = lib().bvo_read_offsets(C.byref(p), offs.ctypes.data, offs.size, off.ctypes.data)
    if rc:
        raise _err(rc)
    part = np.ascontiguousarray(degrees[lo:hi])
    out = np.zeros(int(part.sum()), dtype=np.int64)
    got = lib().bvo_decode_labels(kind, width, lab.ctypes.data, lab.size, off.ctypes.data, lo, hi, part.ctypes.data, out.ctypes.data)
    if got != out.size:
        raise ValueError("label stream and label offsets disagree")
    return out
