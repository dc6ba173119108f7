"""The input generator / encoder (datagen/): bit-exact with the reference compressor on the fixture, and
round-trips through the independent oracle for every coding and parameter combination."""
import itertools

import numpy as np
import pytest

import bvgen as bg
import bvoracle as bo


def oracle_of(e):
    p = bo.params_from_properties(bo.parse_properties(e.properties_text()))
    return bo.OracleGraph(p, e.graph, e.offsets_stream.tobytes())


def test_reencoding_cnr2000_reproduces_the_fixture_bytes(cnr_golden, cnr_basename):
    deg, rowptr, col = cnr_golden
    e = bg.encode_csr(rowptr, col, bg.make_params(window=7, maxref=3, minint=3, zetak=3), nsplits=1, threads=1)
    assert e.graph_bits == 11443904
    assert np.array_equal(e.graph, np.fromfile(cnr_basename + ".graph", dtype=np.uint8))
    assert np.array_equal(e.offsets_stream, np.fromfile(cnr_basename + ".offsets", dtype=np.uint8))


@pytest.mark.parametrize("coding,k", [(bg.GAMMA, 0), (bg.DELTA, 0), (bg.UNARY, 0), (bg.ZETA, 1), (bg.ZETA, 2), (bg.ZETA, 3),
                                      (bg.ZETA, 5), (bg.NIBBLE, 0), (bg.GOLOMB, 1), (bg.GOLOMB, 3), (bg.GOLOMB, 8)])
def test_every_code_round_trips_through_the_oracle_reader(coding, k):
    rng = np.random.default_rng(coding * 10 + k)
    small = np.arange(0, 70, dtype=np.uint64)
    vals = np.concatenate([small, rng.integers(0, 1 << 20, 300, dtype=np.uint64)])
    if coding in (bg.GAMMA, bg.DELTA):
        vals = np.concatenate([vals, rng.integers(0, 1 << 62, 200, dtype=np.uint64), np.array([(1 << 63) - 2], dtype=np.uint64)])
    elif coding in (bg.ZETA, bg.NIBBLE):
        vals = np.concatenate([vals, rng.integers(0, 1 << 56, 200, dtype=np.uint64)])
    if coding in (bg.UNARY, bg.GOLOMB):
        vals = vals % np.uint64(3000)
    buf, bits = bg.write_codes(coding, vals, k)
    pos = 0
    for v in vals:
        got, pos = bo.read_code(buf.tobytes() + b"\0" * 16, pos, coding, k)
        assert got == int(v)
    assert pos == bits


def tree_csr(n, inward):
    """Complete binary trees as in BVGraphTest.testCompression (heap numbering: children of i are 2i+1, 2i+2)."""
    rows = [[] for _ in range(n)]
    for i in range(n):
        for c in (2 * i + 1, 2 * i + 2):
            if c < n:
                (rows[c] if inward else rows[i]).append(i if inward else c)
    rowptr = np.zeros(n + 1, dtype=np.int64)
    rowptr[1:] = np.cumsum([len(r) for r in rows])
    col = np.array([v for r in rows for v in sorted(r)], dtype=np.int64)
    return rowptr, col


def test_tree_grid_round_trip():
    """The w x r x i grid of BVGraphTest.java:52-103 (flags 0), decoded by the oracle."""
    for n, w, r, i, inward in itertools.product(range(1, 8), range(3), range(3), range(4), (False, True)):
        rowptr, col = tree_csr(n, inward)
        e = bg.encode_csr(rowptr, col, bg.make_params(window=w, maxref=r, minint=i))
        d, c = oracle_of(e).decode_range()
        assert np.array_equal(d, np.diff(rowptr)) and np.array_equal(c, col), (n, w, r, i, inward)
        assert e.arcs == col.size


@pytest.mark.parametrize("residual", [bg.ZETA, bg.GAMMA, bg.DELTA, bg.NIBBLE, bg.GOLOMB])
@pytest.mark.parametrize("other", [bg.GAMMA, bg.DELTA])
def test_generated_graphs_round_trip_with_every_flag(residual, other):
    params = bg.make_params(window=5, maxref=4, minint=2, zetak=4, outdegree=other, block=other, residual=residual,
                            reference=bg.UNARY if other == bg.GAMMA else other, blockcount=other, offset=other)
    e = bg.copy_model(3000, seed=11, mean_degree=12, params=params, nsplits=3, threads=2)
    og = oracle_of(e)
    assert np.array_equal(og.offsets, e.offsets)
    d, c = og.decode_range()
    rowptr = np.zeros(d.size + 1, dtype=np.int64)
    np.cumsum(d, out=rowptr[1:])
    assert int(d.sum()) == e.arcs
    assert bg.csr_checksum(0, rowptr, c) == e.checksum
    for x in (0, 1, 17, 999, 1000, 1001, 2999):
        assert np.array_equal(og.successors(x), c[rowptr[x]:rowptr[x + 1]])


def test_output_does_not_depend_on_thread_count():
    a = bg.copy_model(20000, seed=3, nsplits=16, threads=1)
    b = bg.copy_model(20000, seed=3, nsplits=16, threads=5)
    assert np.array_equal(a.graph, b.graph) and np.array_equal(a.offsets, b.offsets) and a.checksum == b.checksum


def test_unbounded_maxref_builds_long_chains():
    e = bg.copy_model(50000, seed=5, params=bg.make_params(maxref=2147483647), nsplits=1, threads=1)
    d, c = oracle_of(e).decode_range()
    assert int(d.sum()) == e.arcs


def test_big_graph_shape():
    e = bg.big_graph(5000, giant_every=0, giant_degree=1000, giant_stride=4, nsplits=2)
    d, c = oracle_of(e).decode_range()
    assert d[0] == 1000 and (d[1:] == 2).all()
    assert np.array_equal(c[:1000], np.arange(1000) * 4)
