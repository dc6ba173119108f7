"""Pins the CPU oracle on the reference's own known-answer fixture cnr-2000 (the assertions of
test/it/unimi/dsi/big/webgraph/BVGraphTest.java:105-123 and WebGraphTestCase.java:106-199 of the reference)."""
import numpy as np
import pytest

import bvoracle as bo


@pytest.fixture(scope="module")
def og(cnr_basename):
    return bo.OracleGraph.load(cnr_basename)


def test_properties(og):
    p = og.p
    assert (p.nodes, p.arcs, p.window, p.maxref, p.minint, p.zetak) == (325557, 3216152, 7, 3, 3, 3)
    assert (p.outdegree_coding, p.block_coding, p.residual_coding, p.reference_coding, p.blockcount_coding, p.offset_coding) == \
        (bo.GAMMA, bo.GAMMA, bo.ZETA, bo.UNARY, bo.GAMMA, bo.GAMMA)


def test_offsets_land_on_the_stream_length(og):
    assert og.offsets[0] == 0
    assert og.offsets[-1] == 11443904 <= og.graph.size * 8
    assert (np.diff(og.offsets.astype(np.int64)) > 0).all()


def test_sequential_scan_equals_golden(og, cnr_golden):
    deg, rowptr, col = cnr_golden
    d, c = og.decode_range()
    assert np.array_equal(d, deg)
    assert np.array_equal(c, col)
    assert int(d.sum()) == og.p.arcs  # WebGraphTestCase.java:140


def test_split_scan_equals_golden(og, cnr_golden):
    deg, rowptr, col = cnr_golden
    for k in (2, 7):
        d, c = og.decode_range(threads=k)
        assert np.array_equal(d, deg) and np.array_equal(c, col)


def test_random_access_equals_golden(og, cnr_golden):
    deg, rowptr, col = cnr_golden
    rng = np.random.default_rng(7)
    for x in np.concatenate([[0, 1, 7, 8, og.p.nodes - 1], rng.integers(0, og.p.nodes, 5000)]):
        x = int(x)
        assert og.outdegree(x) == deg[x]
        assert np.array_equal(og.successors(x), col[rowptr[x]:rowptr[x + 1]])


def test_node_iterator_from_every_kind_of_start(og, cnr_golden):
    """nodeIterator(s) agrees with random access (WebGraphTestCase.java:151-180): exercises the window pre-fill."""
    deg, rowptr, col = cnr_golden
    n = og.p.nodes
    for s in (1, 2, 6, 7, 8, 9, 1000, 123457, n - 3, n):
        e = min(s + 2000, n)
        d, c = og.decode_range(s, e)
        assert np.array_equal(d, deg[s:e])
        assert np.array_equal(c, col[rowptr[s]:rowptr[e]])


def test_index_errors(og):
    with pytest.raises(IndexError):
        og.successors(og.p.nodes)
    with pytest.raises(IndexError):
        og.outdegree(-1)


def test_flag_parsing():
    assert bo.string2flags("") == 0
    f = bo.string2flags("RESIDUALS_DELTA | OUTDEGREES_DELTA|REFERENCES_GAMMA")
    assert f == (bo.DELTA << 8) | bo.DELTA | (bo.GAMMA << 12)
    with pytest.raises(IOError):
        bo.string2flags("RESIDUALS_UNARY")
    with pytest.raises(IOError):
        bo.params_from_properties({"graphclass": "it.unimi.dsi.big.webgraph.EFGraph", "version": "0"})
    with pytest.raises(IOError):
        bo.params_from_properties({"graphclass": "it.unimi.dsi.webgraph.BVGraph"})


def test_transpose_restatement(cnr_golden):
    """oracle.transpose_csr (Transform.transpose restated): a hand-checked case, arc conservation and the involution the
    reference's own TransformTest relies on, on the golden fixture."""
    import bvoracle as bo
    rp = np.array([0, 2, 3, 3, 5], dtype=np.int64)
    col = np.array([1, 3, 0, 0, 2], dtype=np.int64)
    trp, tcol = bo.transpose_csr(4, rp, col)
    assert trp.tolist() == [0, 2, 3, 4, 5] and tcol.tolist() == [1, 3, 0, 3, 0]
    deg, rowptr, gcol = cnr_golden
    n = deg.size
    trp, tcol = bo.transpose_csr(n, rowptr, gcol)
    assert trp[-1] == rowptr[-1] and np.all(np.diff(trp) >= 0)
    brp, bcol = bo.transpose_csr(n, trp, tcol)
    assert np.array_equal(brp, rowptr) and np.array_equal(bcol, gcol)


def test_label_streams_round_trip():
    """datagen.write_labels (BitStreamArcLabelledImmutableGraph.store restated) against oracle.decode_labels (its node
    iterator restated): fixed widths 0..32 and gamma-coded labels, whole graph and a sub-range, and a hand-checked
    stream (three 3-bit labels 5, 0, 7 on a node of degree 3 = bits 101 000 111)."""
    import bvgen as bg
    import bvoracle as bo
    lab, off = bg.write_labels(bg.LABEL_FIXED_INT, 3, np.array([0, 3, 3]), np.array([5, 0, 7]))
    assert lab.tolist() == [0b10100011, 0b10000000]
    assert off.tolist() == [0b10001010, 0b10000000]  # gamma(0) = 1, gamma(9) = 0001010, gamma(0) = 1
    rng = np.random.default_rng(5)
    deg = rng.integers(0, 12, 2000)
    rp = np.zeros(2001, dtype=np.int64)
    np.cumsum(deg, out=rp[1:])
    for kind, width, hi in ((0, 0, 1), (0, 1, 2), (0, 7, 128), (0, 31, 1 << 31), (0, 32, 1 << 32), (1, 0, 1 << 24)):
        vals = rng.integers(0, hi, rp[-1])
        lab, off = bg.write_labels(kind, width, rp, vals)
        got = bo.decode_labels(kind, width, lab, off, deg)
        assert np.array_equal(got.astype(np.uint32), vals.astype(np.uint32))
        assert np.array_equal(bo.decode_labels(kind, width, lab, off, deg, 500, 1500), got[rp[500]:rp[1500]])
    # the benchmark's labels: arc j carries the low `width` bits of (j * multiplier mod 2^64) >> 20
    lab, off = bg.write_labels_synthetic(10, rp)
    j = np.arange(rp[-1], dtype=np.uint64)
    expect = ((j * np.uint64(bg.LABEL_HASH_MULTIPLIER)) >> np.uint64(20)) & np.uint64(1023)
    assert np.array_equal(bo.decode_labels(0, 10, lab, off, deg).astype(np.uint64), expect)


def test_label_list_streams_round_trip():
    """List-valued labels: datagen.write_label_lists (FixedWidthIntListLabel.toBitStream restated) against
    oracle.decode_label_lists (fromBitStream restated), and a hand-checked stream: one arc with the list [5, 2] of 3-bit
    elements = gamma(2) 011, then 101 010."""
    import bvgen as bg
    lab, off = bg.write_label_lists(3, np.array([0, 1]), np.array([0, 2]), np.array([5, 2]))
    assert lab.tolist() == [0b01110101, 0b00000000]
    lp, vals = bo.decode_label_lists(bo.LABEL_FIXED_INT_LIST, 3, lab, off, np.array([1]))
    assert lp.tolist() == [0, 2] and vals.tolist() == [5, 2]
    rng = np.random.default_rng(4)
    deg = rng.integers(0, 7, 1500)
    rp = np.concatenate([[0], np.cumsum(deg)]).astype(np.int64)
    for kind, width in ((bo.LABEL_FIXED_INT_LIST, 0), (bo.LABEL_FIXED_INT_LIST, 13), (bo.LABEL_FIXED_INT_LIST, 32), (bo.LABEL_FIXED_LONG_LIST, 47),
                        (bo.LABEL_FIXED_LONG_LIST, 64)):
        lens = rng.integers(0, 9, int(rp[-1]))
        lp = np.concatenate([[0], np.cumsum(lens)]).astype(np.int64)
        raw = rng.integers(0, 1 << 63, int(lp[-1]), dtype=np.uint64) * 2 + 1
        values = (raw >> np.uint64(64 - width) if width else raw * np.uint64(0)).view(np.int64)
        lab, off = bg.write_label_lists(width, rp, lp, values)
        glp, gv = bo.decode_label_lists(kind, width, lab, off, deg)
        want = values if kind == bo.LABEL_FIXED_LONG_LIST else values.astype(np.uint32).view(np.int32).astype(np.int64)
        assert np.array_equal(glp, lp) and np.array_equal(gv, want)
        glp, gv = bo.decode_label_lists(kind, width, lab, off, deg, 500, 900)
        assert np.array_equal(glp, lp[rp[500]:rp[900] + 1] - lp[rp[500]]) and np.array_equal(gv, want[lp[rp[500]]:lp[rp[900]]])
