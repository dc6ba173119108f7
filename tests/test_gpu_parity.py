"""GPU parity tests (run with -m gpu on a B200): the CUDA decode, called through the C ABI, against the reference's
golden fixture and against the CPU oracle on the same seeded inputs.  Bit-exact: everything is integer work."""
import numpy as np
import pytest

import bvgen as bg
import bvoracle as bo
import webgraph_big_b200 as wg

pytestmark = pytest.mark.gpu


def oracle_of(e):
    p = bo.params_from_properties(bo.parse_properties(e.properties_text()))
    return bo.OracleGraph(p, e.graph, e.offsets_stream.tobytes())


def gpu_of(e, **kw):
    return wg.GpuBVGraph.load_mem(e.properties_text(), e.graph, e.offsets_stream, **kw)


def rowptr_of(deg):
    rp = np.zeros(deg.size + 1, dtype=np.int64)
    np.cumsum(deg, out=rp[1:])
    return rp


# ------------------------------------------------------------------------------------------------ the golden fixture
@pytest.fixture(scope="module")
def cnr(cnr_basename):
    g = wg.GpuBVGraph.load(cnr_basename)
    yield g
    g.close()


def test_cnr2000_bulk_decode_equals_golden(cnr, cnr_golden):
    """BVGraphTest.testLarge (BVGraphTest.java:105-123): the loaded graph equals the ASCII golden."""
    deg, rowptr, col = cnr_golden
    assert cnr.numNodes() == 325557 and cnr.numArcs() == 3216152
    rp, c = cnr.decodeToCSR()
    assert np.array_equal(rp, rowptr)
    assert np.array_equal(c, col)


def test_cnr2000_random_access_equals_golden(cnr, cnr_golden):
    """BVGraphTest.java:114-121: outdegree(i) and successors(i) for every node, including the trailing -1."""
    deg, rowptr, col = cnr_golden
    ids = np.arange(cnr.numNodes(), dtype=np.int64)
    assert np.array_equal(cnr.outdegrees(ids), deg)
    rp, c = cnr.successorsBatch(ids)
    assert np.array_equal(rp, rowptr) and np.array_equal(c, col)
    rng = np.random.default_rng(3)
    ids = rng.integers(0, cnr.numNodes(), 20000)
    rp, c = cnr.successorsBatch(ids)
    want = np.concatenate([col[rowptr[x]:rowptr[x + 1]] for x in ids])
    assert np.array_equal(np.diff(rp), deg[ids]) and np.array_equal(c, want)
    it = cnr.successors(5)
    got = []
    while (v := it.nextLong()) != -1:
        got.append(v)
    assert got == list(col[rowptr[5]:rowptr[6]]) and it.nextLong() == -1


def test_cnr2000_node_iterator_from_any_start(cnr, cnr_golden):
    """WebGraphTestCase.assertGraph (WebGraphTestCase.java:151-180): nodeIterator(s) agrees with random access."""
    deg, rowptr, col = cnr_golden
    n = cnr.numNodes()
    for s in (0, 1, 6, 7, 8, 22, 1000, 123457, n - 3, n):
        e = min(s + 3000, n)
        rp, c = cnr.decodeToCSR(s, e)
        assert np.array_equal(rp, rowptr[s:e + 1] - rowptr[s]), s
        assert np.array_equal(c, col[rowptr[s]:rowptr[e]]), s
    it = cnr.nodeIterator(n - 40)
    with pytest.raises(wg.IllegalStateException):  # BVG:1185
        it.outdegree()
    seen = 0
    while it.hasNext():
        x = it.nextLong()
        assert it.outdegree() == deg[x]
        assert np.array_equal(it.successorArray(), col[rowptr[x]:rowptr[x + 1]])
        seen += 1
    assert seen == 40
    with pytest.raises(wg.NoSuchElementException):  # BVG:1165
        it.nextLong()


def test_cnr2000_split_iterators_and_copies(cnr, cnr_golden):
    """splitNodeIterators (IG:405-436) over flyweight copies: the parts cover the graph exactly once."""
    deg, rowptr, col = cnr_golden
    parts = cnr.splitNodeIterators(5)
    assert len(parts) == 5
    total = 0
    nxt = 0
    for it in parts:
        while it.hasNext():
            x = it.nextLong()
            assert x == nxt
            nxt += 1
            total += it.outdegree()
            if x % 997 == 0:
                assert np.array_equal(it.successorArray(), col[rowptr[x]:rowptr[x + 1]])
    assert nxt == cnr.numNodes() and total == cnr.numArcs()
    c = cnr.copy()
    assert np.array_equal(c.successorArray(12345), col[rowptr[12345]:rowptr[12346]])
    c.close()


def test_index_errors(cnr):
    n = cnr.numNodes()
    for call in (lambda: cnr.outdegree(n), lambda: cnr.successors(-1), lambda: cnr.successorArray(n), lambda: cnr.nodeIterator(n + 1),
                 lambda: cnr.decodeToCSR(5, n + 1), lambda: cnr.successorsBatch(np.array([1, n]))):
        with pytest.raises(wg.IllegalArgumentException):
            call()
    rp, c = cnr.decodeToCSR(7, 7)
    assert rp.tolist() == [0] and c.size == 0


def test_empty_range_on_a_new_handle_and_upload_exclusivity(cnr_basename, cnr_golden):
    """decodeToCSRDevice(x, x) as the first call of a handle (nothing has sized the stream's scratch yet) leaves
    rowptr[0] = 0 and touches nothing else; re-uploading the bit stream is refused while a second stream shares the graph."""
    torch = pytest.importorskip("torch")
    deg, rowptr, col = cnr_golden
    with wg.GpuBVGraph.load(cnr_basename) as g:
        d_rp = torch.full((4,), -7, dtype=torch.int64, device="cuda")
        d_col = torch.full((4,), -7, dtype=torch.int64, device="cuda")
        for x in (0, 7, g.numNodes()):
            g.decodeToCSRDevice(x, x, d_rp.data_ptr(), d_col.data_ptr(), 4)
            g.sync()
            assert d_rp.tolist() == [0, -7, -7, -7] and d_col.tolist() == [-7] * 4
        rp, c = g.decodeToCSR(0, 100)
        assert np.array_equal(c, col[:rowptr[100]])
        raw = np.fromfile(cnr_basename + ".graph", dtype=np.uint8)
        it = g.scan(0, 1000, 500, upload=(raw.ctypes.data, raw.size))
        assert sum(cc.size for _, _, _, cc in it) == rowptr[1000]
        other = g.copy()
        with pytest.raises(wg.IllegalStateException):
            list(g.scan(0, 1000, 500, upload=(raw.ctypes.data, raw.size)))
        other.close()


def test_device_range_decodes_with_halo_and_exact_capacity():
    """wgbv_decode_csr (device buffers) on ranges that start in the middle of a compression split: the front end runs
    inside k_extras after the arcs of the halo rows have been counted.  Every range is decoded into a buffer of exactly
    its size with a guard behind it; a buffer one element short is refused (IllegalArgumentException, the reference's
    answer to a too-short array) and nothing is written behind the capacity given.  The graph has rows of hundreds of
    successors, so the capacity checks of the tile kernel and of the heavy-row kernel both run."""
    torch = pytest.importorskip("torch")
    e = bg.copy_model(60000, seed=41, mean_degree=30, max_degree=3000, nsplits=3)
    deg, col = oracle_of(e).decode_range()
    rowptr = rowptr_of(deg)
    n = deg.size
    GUARD = -1234567
    with gpu_of(e) as g:
        rng = np.random.default_rng(8)
        cuts = [(1, n), (7, 8), (20, 41), (1023, 1025), (1024, 5000), (19999, 20001), (20000, 40000), (20021, 20022), (n - 1, n)]
        cuts += [tuple(sorted(rng.integers(0, n + 1, 2).tolist())) for _ in range(12)]
        for a, b in cuts:
            if a == b:
                continue
            arcs = int(rowptr[b] - rowptr[a])
            d_rp = torch.full((b - a + 1 + 8,), GUARD, dtype=torch.int64, device="cuda")
            d_col = torch.full((arcs + 64,), GUARD, dtype=torch.int64, device="cuda")
            g.decodeToCSRDevice(a, b, d_rp.data_ptr(), d_col.data_ptr(), arcs)
            rp, c = d_rp.cpu().numpy(), d_col.cpu().numpy()
            assert np.array_equal(rp[: b - a + 1], rowptr[a:b + 1] - rowptr[a]) and np.all(rp[b - a + 1:] == GUARD), (a, b)
            assert np.array_equal(c[:arcs], col[rowptr[a]:rowptr[b]]) and np.all(c[arcs:] == GUARD), (a, b)
            if arcs > 1:
                short = arcs - 1 if arcs < 50 else arcs - int(rng.integers(1, min(arcs, 4000)))
                d_col.fill_(GUARD)
                with pytest.raises(wg.IllegalArgumentException):
                    g.decodeToCSRDevice(a, b, d_rp.data_ptr(), d_col.data_ptr(), short)
                assert bool((d_col[short:] == GUARD).all()), (a, b, short)
        # and the handle still works
        d_rp = torch.empty(n + 1, dtype=torch.int64, device="cuda")
        d_col = torch.empty(col.size, dtype=torch.int64, device="cuda")
        g.decodeToCSRDevice(0, n, d_rp.data_ptr(), d_col.data_ptr(), col.size)
        assert np.array_equal(d_col.cpu().numpy(), col)


def test_graph_without_offsets_file(cnr_basename, cnr_golden):
    """A graph shipped without its .offsets (the reference needs `BVGraph -o` first, BVG:2595-2609): wgbv_open_mem with
    offsets == NULL finds the records itself and the decode equals the golden."""
    deg, rowptr, col = cnr_golden
    props = open(cnr_basename + ".properties").read()
    graph = np.fromfile(cnr_basename + ".graph", dtype=np.uint8)
    with wg.GpuBVGraph.load_mem(props, graph, None) as g:
        rp, c = g.decodeToCSR()
        assert np.array_equal(rp, rowptr) and np.array_equal(c, col)
        ids = np.array([0, 17, 325556, 100000], dtype=np.int64)
        _, c2 = g.successorsBatch(ids)
        assert np.array_equal(c2, np.concatenate([col[rowptr[x]:rowptr[x + 1]] for x in ids]))


def test_properties_that_understate_the_arc_count():
    """The scratch list of long rows is sized from the arc count the .properties promise.  A file that holds far more arcs
    than it says (here: 1000 instead of millions, with thousands of rows of hundreds of successors) is reported
    (IllegalStateException) instead of writing past the list; the same graph with honest properties decodes."""
    import re
    torch = pytest.importorskip("torch")
    e = bg.copy_model(30000, seed=6, mean_degree=300, max_degree=20000, nsplits=2)
    deg, col = oracle_of(e).decode_range()
    lying = re.sub(r"(?m)^arcs=\d+$", "arcs=1000", e.properties_text())
    assert lying != e.properties_text()
    d_rp = torch.empty(e.n + 1, dtype=torch.int64, device="cuda")
    d_col = torch.empty(col.size, dtype=torch.int64, device="cuda")
    with wg.GpuBVGraph.load_mem(lying, e.graph, e.offsets_stream) as g:
        with pytest.raises(wg.IllegalStateException):
            g.decodeToCSRDevice(0, e.n, d_rp.data_ptr(), d_col.data_ptr(), col.size)
    with gpu_of(e) as g:
        g.decodeToCSRDevice(0, e.n, d_rp.data_ptr(), d_col.data_ptr(), col.size)
        assert np.array_equal(d_col.cpu().numpy(), col)


def test_checksum_kernel_matches_generator_hash(cnr, cnr_golden):
    torch = pytest.importorskip("torch")
    deg, rowptr, col = cnr_golden
    n = cnr.numNodes()
    d_rp = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    arcs = cnr.decodeDegreesDevice(0, n, d_rp.data_ptr())
    assert arcs == col.size
    d_col = torch.empty(arcs, dtype=torch.int64, device="cuda")
    cnr.decodeToCSRDevice(0, n, d_rp.data_ptr(), d_col.data_ptr(), arcs)
    assert np.array_equal(d_col.cpu().numpy(), col)
    assert cnr.csrChecksumDevice(0, n, d_rp.data_ptr(), d_col.data_ptr()) == bg.csr_checksum(0, rowptr, col)
    ms = cnr.lastStageMs()
    assert len(ms) == 5 and ms[4] > 0
    assert cnr.kernelLaunches() > 0
    assert cnr.algorithmicBytes(0, n, arcs) == (11443904 + 7) // 8 + (n * (2 + 5) + 7) // 8 + 8 * (n + 1) + 8 * arcs


# ------------------------------------------------------------------------------------------------ synthetic inputs vs oracle
@pytest.mark.parametrize("window,maxref,minint", [(7, 3, 4), (0, 3, 4), (1, 1, 0), (3, 0, 2), (16, 8, 1), (7, 3, 3)])
def test_copy_model_equals_oracle(window, maxref, minint):
    e = bg.copy_model(40000, seed=window * 100 + maxref * 10 + minint, mean_degree=15,
                      params=bg.make_params(window=window, maxref=maxref, minint=minint), nsplits=4)
    deg, col = oracle_of(e).decode_range()
    with gpu_of(e) as g:
        rp, c = g.decodeToCSR()
        assert np.array_equal(rp, rowptr_of(deg)) and np.array_equal(c, col)
        for s, t in ((1, 40000), (39990, 40000), (20000, 20001), (12345, 23456)):
            rp2, c2 = g.decodeToCSR(s, t)
            assert np.array_equal(c2, col[rp[s]:rp[t]]) and np.array_equal(rp2, rp[s:t + 1] - rp[s])
        ids = np.random.default_rng(1).integers(0, 40000, 3000)
        rpb, cb = g.successorsBatch(ids)
        assert np.array_equal(cb, np.concatenate([col[rp[x]:rp[x + 1]] for x in ids]))


@pytest.mark.parametrize("residual", [bg.ZETA, bg.GAMMA, bg.DELTA, bg.NIBBLE, bg.GOLOMB])
@pytest.mark.parametrize("other", [bg.GAMMA, bg.DELTA])
@pytest.mark.parametrize("zetak", [1, 3, 5])
def test_every_compression_flag_equals_oracle(residual, other, zetak):
    """Codings the reference's own tests never exercise (parity with the reference itself is unpinned for them;
    here they are checked against the oracle, which reads them from the published definitions)."""
    params = bg.make_params(window=5, maxref=4, minint=2, zetak=zetak, outdegree=other, block=other, residual=residual,
                            reference=bg.UNARY if other == bg.GAMMA else other, blockcount=other, offset=other)
    e = bg.copy_model(6000, seed=31 + residual + other, mean_degree=12, params=params, nsplits=3)
    deg, col = oracle_of(e).decode_range()
    with gpu_of(e) as g:
        rp, c = g.decodeToCSR()
        assert np.array_equal(rp, rowptr_of(deg)) and np.array_equal(c, col)


def test_erdos_renyi_config1():
    """BASELINE config 1 at reduced size: ER, w=7, maxRef=3, minInt=4, zeta3 (residual-dominated rows)."""
    e = bg.erdos_renyi(100000, 1.6e-4, seed=1, params=bg.make_params(7, 3, 4, 3))
    deg, col = oracle_of(e).decode_range()
    with gpu_of(e) as g:
        rp, c = g.decodeToCSR()
        assert np.array_equal(rp, rowptr_of(deg)) and np.array_equal(c, col)


def test_unbounded_reference_chains_config3():
    """maxrefcount=2147483647 (what -m -1 writes, BVG:2654): deep level-by-level copy resolution."""
    e = bg.copy_model(60000, seed=9, mean_degree=10, p_ref=0.95, params=bg.make_params(maxref=2147483647), nsplits=1)
    deg, col = oracle_of(e).decode_range()
    with gpu_of(e) as g:
        rp, c = g.decodeToCSR()
        assert np.array_equal(rp, rowptr_of(deg)) and np.array_equal(c, col)
        rp2, c2 = g.decodeToCSR(50000, 60000)
        assert np.array_equal(c2, col[rp[50000]:rp[60000]])
        ids = np.array([59999, 1, 31234, 44444], dtype=np.int64)
        rpb, cb = g.successorsBatch(ids)
        assert np.array_equal(cb, np.concatenate([col[rp[x]:rp[x + 1]] for x in ids]))


def test_giant_rows_and_zero_degree_nodes():
    """The shape of the reference's slow test (slow/.../BVGraphSlowTest.java:30-85): a row far larger than all others."""
    e = bg.big_graph(30000, giant_every=7000, giant_degree=200000, giant_stride=3, nsplits=2)
    deg, col = oracle_of(e).decode_range()
    with gpu_of(e) as g:
        rp, c = g.decodeToCSR()
        assert np.array_equal(rp, rowptr_of(deg)) and np.array_equal(c, col)
    # isolated nodes only
    rowptr = np.zeros(1001, dtype=np.int64)
    e = bg.encode_csr(rowptr, np.zeros(0, dtype=np.int64))
    with gpu_of(e) as g:
        rp, c = g.decodeToCSR()
        assert not rp.any() and c.size == 0


@pytest.mark.gpu
@pytest.mark.parametrize("fused", ["0", "1"])
def test_giant_rows_in_a_web_like_graph(monkeypatch, fused):
    """BASELINE config 5's shape in small: a sparse copy-model graph with a few rows of tens of thousands of successors
    (residuals for stride 4, one long interval for stride 1; other rows reference them), decoded by a whole warp each."""
    monkeypatch.setenv("WGBV_FUSED", fused)
    for stride, gd in ((4, 60000), (1, 150000), (7, 30000)):
        e = bg.copy_model(250000, seed=50 + stride, mean_degree=3.3, nsplits=4, giant_every=100000, giant_degree=gd, giant_stride=stride)
        deg, col = oracle_of(e).decode_range()
        assert deg.max() >= 30000
        with gpu_of(e) as g:
            rp, c = g.decodeToCSR()
            assert np.array_equal(rp, rowptr_of(deg)) and np.array_equal(c, col)
            rp, c = g.decodeToCSR(49990, 150010)
            assert np.array_equal(c, col[rowptr_of(deg)[49990]:rowptr_of(deg)[150010]])


@pytest.mark.parametrize("fused", ["0", "1"])
def test_repeated_and_concurrent_decodes_are_identical(monkeypatch, cnr_basename, cnr_golden, fused):
    """The kernels hand work out by tickets and pass prefix sums, level lists and row data between blocks through global
    memory (decoupled look-back, in-place merges).  compute-sanitizer is closed on the pool these tests run on, so the
    protocols are stressed instead: the same decodes many times over, on two streams of one graph at the same time and on
    ranges that cut tiles at odd places -- every result must equal the golden, bit for bit, every time."""
    import threading
    monkeypatch.setenv("WGBV_FUSED", fused)
    deg, rowptr, col = cnr_golden
    n = deg.size
    e = bg.copy_model(120000, seed=4242, mean_degree=18, nsplits=3)
    odeg, ocol = oracle_of(e).decode_range()
    orp = rowptr_of(odeg)
    errors = []

    def work(g, want_rp, want_col, ranges, reps):
        try:
            for k in range(reps):
                for lo, hi in ranges:
                    rp, c = g.decodeToCSR(lo, hi)
                    if not (np.array_equal(rp, want_rp[lo:hi + 1] - want_rp[lo]) and np.array_equal(c, want_col[want_rp[lo]:want_rp[hi]])):
                        errors.append((k, lo, hi))
        except Exception as ex:  # noqa: BLE001
            errors.append(repr(ex))

    with wg.GpuBVGraph.load(cnr_basename) as g1, gpu_of(e) as g2:
        c1, c2 = g1.copy(), g2.copy()
        ths = [threading.Thread(target=work, args=(g1, rowptr, col, [(0, n), (1017, 250001)], 8)),
               threading.Thread(target=work, args=(c1, rowptr, col, [(300000, n), (0, 99999), (5, 4101)], 8)),
               threading.Thread(target=work, args=(g2, orp, ocol, [(0, e.n), (33, 77777)], 10)),
               threading.Thread(target=work, args=(c2, orp, ocol, [(100000, e.n), (0, 60001)], 10))]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        c1.close()
        c2.close()
    assert not errors, errors[:5]


def test_tree_grid():
    """BVGraphTest.testCompression grid (BVGraphTest.java:52-103), a sample of it."""
    from test_datagen import tree_csr
    for n, w, r, i, inward in ((1, 0, 0, 0, False), (7, 2, 2, 3, True), (7, 1, 0, 1, False), (5, 2, 1, 2, True), (3, 0, 2, 0, False)):
        rowptr, col = tree_csr(n, inward)
        e = bg.encode_csr(rowptr, col, bg.make_params(window=w, maxref=r, minint=i))
        with gpu_of(e) as g:
            rp, c = g.decodeToCSR()
            assert np.array_equal(rp, rowptr) and np.array_equal(c, col)
            for s in range(n + 1):
                rp2, c2 = g.decodeToCSR(s, n)
                assert np.array_equal(c2, col[rowptr[s]:])


def test_shards_concatenate_to_the_whole_graph():
    """Bit-balanced contiguous shards with a window*maxref halo decode independently; the CSR is their concatenation."""
    e = bg.copy_model(50000, seed=77, mean_degree=15, nsplits=8)
    deg, col = oracle_of(e).decode_range()
    rowptr = rowptr_of(deg)
    for count in (2, 3, 8):
        nxt, total, chk = 0, 0, 0
        for k in range(count):
            with gpu_of(e, shard_index=k, shard_count=count) as g:
                lo, hi = g.shardRange()
                assert lo == nxt
                nxt = hi
                rp, c = g.decodeToCSR()
                assert np.array_equal(rp, rowptr[lo:hi + 1] - rowptr[lo]) and np.array_equal(c, col[rowptr[lo]:rowptr[hi]])
                total += c.size
                chk = (chk + bg.csr_checksum(lo, rp, c)) % (1 << 64)
                if hi > lo:
                    with pytest.raises(wg.IllegalArgumentException):
                        g.decodeToCSR(lo, hi + 1) if hi < 50000 else g.decodeToCSR(lo - 1, hi) if lo > 0 else g.decodeToCSR(-1, hi)
        assert nxt == 50000 and total == e.arcs and chk == e.checksum
    with pytest.raises(wg.UnsupportedOperationException):
        e2 = bg.copy_model(1000, seed=1, params=bg.make_params(maxref=2147483647), nsplits=1)
        gpu_of(e2, shard_index=0, shard_count=2)


def test_malformed_reference_is_reported():
    """A reference larger than the window is the reference's IllegalStateException (BVG:701)."""
    e = bg.copy_model(2000, seed=5, params=bg.make_params(window=7), nsplits=1)
    text = e.properties_text().replace("windowsize=7", "windowsize=1")
    g = wg.GpuBVGraph.load_mem(text, e.graph, e.offsets_stream)
    with pytest.raises(wg.IllegalStateException):
        g.decodeToCSR()
    g.close()


def _similar_rows_graph(n, centers, width, size, seed):
    """Rows around each centre hold large, nearly identical successor lists (so the encoder chains references across
    tile boundaries); everything else is tiny."""
    rng = np.random.default_rng(seed)
    rows = [np.array([x], dtype=np.int64) if x % 3 else np.zeros(0, dtype=np.int64) for x in range(n)]
    for c in centers:
        base = np.unique(rng.integers(0, n, size))
        for x in range(c - width, c + width):
            keep = base[rng.random(base.size) < 0.97]
            rows[x] = np.unique(np.concatenate([keep, rng.integers(0, n, 5)]))
    rowptr = np.zeros(n + 1, dtype=np.int64)
    rowptr[1:] = np.cumsum([r.size for r in rows])
    return rowptr, np.concatenate(rows)


def test_tiles_that_do_not_fit_shared_memory(monkeypatch):
    """The one-pass tile kernel's slow tiles: rows a few thousand (or tens of thousands of) successors long that reference
    each other across tile boundaries are decoded by k_slow in global memory; ranges that start next to them."""
    monkeypatch.setenv("WGBV_FUSED", "1")
    for size, centers in ((4000, (256, 1000, 1536)), (40000, (512, 2048 + 256))):
        n = 3000
        rowptr, col = _similar_rows_graph(n, centers, 12, size, size)
        e = bg.encode_csr(rowptr, col, bg.make_params(window=7, maxref=3, minint=4))
        deg, ocol = oracle_of(e).decode_range()
        assert np.array_equal(ocol, col)
        with gpu_of(e) as g:
            rp, c = g.decodeToCSR()
            assert np.array_equal(rp, rowptr) and np.array_equal(c, col)
            st = g.fusedStats()
            assert st["enabled"] and st["slow_tiles"] > 0 and st["reruns"] == 0
            # ranges that start inside the run of long rows (the first tile is slow and its rows reference rows before the range)
            for lo in (centers[0] + 3, centers[0] - 5, centers[1] + 1):
                rp2, c2 = g.decodeToCSR(lo, n)
                assert np.array_equal(rp2, rowptr[lo:] - rowptr[lo]) and np.array_equal(c2, col[rowptr[lo]:])


def test_fused_and_general_paths_agree(monkeypatch, cnr_basename, cnr_golden):
    """The one-pass tile kernel (default) and the level-synchronous kernels (WGBV_FUSED=0) on the golden fixture, on
    ranges that start inside a tile, and on synthetic graphs with 1..8 levels and rows on both merge paths."""
    deg, rowptr, col = cnr_golden
    graphs = [bg.copy_model(60000, seed=7 + maxref, mean_degree=md, params=bg.make_params(window=window, maxref=maxref, minint=4), nsplits=2)
              for maxref, window, md in ((1, 1, 12), (3, 7, 25), (4, 7, 60), (8, 8, 25))]
    oracles = [oracle_of(e).decode_range() for e in graphs]
    for fused in ("1", "0"):
        monkeypatch.setenv("WGBV_FUSED", fused)
        with wg.GpuBVGraph.load(cnr_basename) as g:
            assert g.fusedStats()["enabled"] == (fused == "1")
            rp, c = g.decodeToCSR()
            assert np.array_equal(rp, rowptr) and np.array_equal(c, col)
            for lo, hi in ((100001, 200003), (1, 2), (325000, 325557), (0, 0), (5, 5)):
                rp, c = g.decodeToCSR(lo, hi)
                assert np.array_equal(rp, rowptr[lo:hi + 1] - rowptr[lo]) and np.array_equal(c, col[rowptr[lo]:rowptr[hi]])
        for e, (odeg, ocol) in zip(graphs, oracles):
            with gpu_of(e) as g:
                rp, c = g.decodeToCSR()
                assert np.array_equal(rp, rowptr_of(odeg)) and np.array_equal(c, ocol)
                assert g.fusedStats()["reruns"] == 0


def test_chains_longer_than_the_properties_promise(monkeypatch):
    """A file whose maxrefcount understates its chains (the reference's decoder never looks at maxrefcount): the tile
    kernel notices, and the decode is rerun on the general path without trusting the property."""
    monkeypatch.setenv("WGBV_FUSED", "1")
    e = bg.copy_model(20000, seed=99, mean_degree=20, params=bg.make_params(window=7, maxref=6, minint=4), nsplits=1)
    odeg, ocol = oracle_of(e).decode_range()
    for lie in (1, 0):
        text = e.properties_text().replace("maxrefcount=6", f"maxrefcount={lie}")
        with wg.GpuBVGraph.load_mem(text, e.graph, e.offsets_stream) as g:
            rp, c = g.decodeToCSR()
            assert np.array_equal(rp, rowptr_of(odeg)) and np.array_equal(c, ocol)
            assert g.fusedStats()["reruns"] >= 1


def test_narrow_scan_delivers_the_same_successors(cnr, cnr_golden):
    """wgbv_scan_set_narrow: successors as uint32, same values, on plain and uploading scans; back to int64 afterwards."""
    deg, rowptr, col = cnr_golden
    n = cnr.numNodes()
    for lo, hi, chunk in ((0, n, 60000), (777, 54321, 5000), (9, 9, 4)):
        got = []
        for a, b, rp, c in cnr.scan(lo, hi, chunk, narrow=True):
            assert c.dtype == np.uint32
            assert np.array_equal(rp, rowptr[a:b + 1] - rowptr[a])
            got.append(c.astype(np.int64))
        assert np.array_equal(np.concatenate(got) if got else np.zeros(0, np.int64), col[rowptr[lo]:rowptr[hi]])
    for a, b, rp, c in cnr.scan(0, 1000, 1000):
        assert c.dtype == np.int64 and np.array_equal(c, col[:rowptr[1000]])


def test_streaming_scan_equals_golden(cnr, cnr_golden):
    """wgbv_scan_*: the pipelined bulk form of nodeIterator(); chunks must tile the range and carry the golden rows."""
    deg, rowptr, col = cnr_golden
    n = cnr.numNodes()
    for lo, hi, chunk in ((0, n, 50000), (12345, 99999, 7777), (5, 6, 10), (n - 1, n, 1), (100, 100, 5)):
        nxt, total = lo, 0
        for a, b, rp, c in cnr.scan(lo, hi, chunk):
            assert a == nxt and b == min(a + chunk, hi)
            nxt = b
            assert np.array_equal(rp, rowptr[a:b + 1] - rowptr[a])
            assert np.array_equal(c, col[rowptr[a]:rowptr[b]])
            total += c.size
        assert nxt == hi and total == rowptr[hi] - rowptr[lo]
    # a scan abandoned half way must not wedge the stream
    it = cnr.scan(0, n, 1000)
    next(it)
    it.close()
    rp, c = cnr.decodeToCSR(0, 100)
    assert np.array_equal(c, col[:rowptr[100]])


def test_streaming_scan_with_upload(cnr_basename, cnr_golden):
    """wgbv_scan_begin_upload: the bit stream is re-sent piece by piece during the scan; results must not change, also
    when the scan is abandoned half way (the rest of the stream is then sent before the handle is used again)."""
    torch = pytest.importorskip("torch")
    deg, rowptr, col = cnr_golden
    raw = np.fromfile(cnr_basename + ".graph", dtype=np.uint8)
    pinned = torch.empty(raw.size + 16, dtype=torch.uint8, pin_memory=True)
    pinned[: raw.size] = torch.from_numpy(raw)
    with wg.GpuBVGraph.load(cnr_basename) as g:
        n = g.numNodes()
        for chunk in (40000, 1 << 20):
            nxt = 0
            for a, b, rp, c in g.scan(0, n, chunk, upload=(pinned.data_ptr(), raw.size)):
                assert a == nxt
                nxt = b
                assert np.array_equal(rp, rowptr[a:b + 1] - rowptr[a]) and np.array_equal(c, col[rowptr[a]:rowptr[b]])
            assert nxt == n
        it = g.scan(0, n, 30000, upload=(pinned.data_ptr(), raw.size))
        next(it)
        it.close()
        rp, c = g.decodeToCSR(n - 5000, n)
        assert np.array_equal(c, col[rowptr[n - 5000]:])


def test_open_slice_with_halo():
    """wgbv_open_slice: a process that holds only its own byte range of the file (plus the window*maxref halo)."""
    e = bg.copy_model(30000, seed=21, mean_degree=14, nsplits=1)
    deg, col = oracle_of(e).decode_range()
    rowptr = rowptr_of(deg)
    lo, hi, halo = 12000, 25000, 21
    first = lo - halo
    byte0 = int(e.offsets[first]) // 8
    byte1 = (int(e.offsets[hi]) + 7) // 8
    rel = e.offsets[first:hi + 1].astype(np.int64) - 8 * byte0
    deltas = np.concatenate([rel[:1], np.diff(rel)]).astype(np.uint64)
    off_stream, _ = bg.write_codes(bg.GAMMA, deltas)
    g = wg.GpuBVGraph.load_slice(e.properties_text(), first, hi - first, e.graph[byte0:byte1], off_stream, halo_nodes=halo)
    assert g.shardRange() == (lo, hi) and g.numNodes() == 30000
    rp, c = g.decodeToCSR()
    assert np.array_equal(rp, rowptr[lo:hi + 1] - rowptr[lo]) and np.array_equal(c, col[rowptr[lo]:rowptr[hi]])
    rp2, c2 = g.successorsBatch(np.array([lo, hi - 1, 20000], dtype=np.int64))
    assert np.array_equal(c2, np.concatenate([col[rowptr[x]:rowptr[x + 1]] for x in (lo, hi - 1, 20000)]))
    with pytest.raises(wg.IllegalArgumentException):
        g.decodeToCSR(lo - 1, hi)
    g.close()


def test_node_ids_beyond_32_bits():
    """A part of a 6*10^9-node graph (the reference's big arrays exist for n > 2^31, BVGraphSlowTest): successors and
    node ids need all 64 bits; the decoded CSR must hash to what the generator hashed before encoding, and random access
    must agree with the bulk decode."""
    first, n = 5_000_000_000, 20000
    e = bg.copy_model(n, seed=5, mean_degree=12, nsplits=1, first_node=first, global_n=6_000_000_000)
    props = e.properties_text().replace(f"nodes={n}", "nodes=6000000000")  # the properties of the whole graph
    g = wg.GpuBVGraph.load_slice(props, first, n, e.graph, e.offsets_stream)
    assert g.numNodes() == 6_000_000_000 and g.shardRange() == (first, first + n)
    rp, c = g.decodeToCSR()
    assert int(rp[-1]) == e.arcs and c.max() > 2 ** 32 and c.min() >= 0
    assert bg.csr_checksum(first, rp, c) == e.checksum
    ids = np.array([first, first + 77, first + n - 1], dtype=np.int64)
    rp2, c2 = g.successorsBatch(ids)
    assert np.array_equal(c2, np.concatenate([c[rp[x - first]:rp[x - first + 1]] for x in ids]))
    with pytest.raises(wg.UnsupportedOperationException):
        next(iter(g.scan(first, first + 10, 10, narrow=True)))
    g.close()


def _transpose_on_gpu(g, n, rowptr, col):
    torch = pytest.importorskip("torch")
    d_rp = torch.from_numpy(np.ascontiguousarray(rowptr, dtype=np.int64)).cuda()
    d_col = torch.from_numpy(np.ascontiguousarray(col, dtype=np.int64)).cuda()
    d_trp = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    d_tcol = torch.empty(max(col.size, 1), dtype=torch.int64, device="cuda")
    g.transposeCSRDevice(n, d_rp.data_ptr(), d_col.data_ptr(), d_trp.data_ptr(), d_tcol.data_ptr())
    torch.cuda.synchronize()
    return d_trp.cpu().numpy(), d_tcol.cpu().numpy()[: col.size]


def test_transpose_equals_oracle_and_is_an_involution(cnr, cnr_golden):
    """wgbv_transpose_csr against the CPU restatement of Transform.transpose, on the golden fixture (two
    radix passes), on its own transpose (the reference's own test: transpose(transpose(g)) == g), and on shapes that
    stress the sort: empty graph, empty rows, a star (one target for every arc), one pass, three passes."""
    deg, rowptr, col = cnr_golden
    n = cnr.numNodes()
    trp, tcol = _transpose_on_gpu(cnr, n, rowptr, col)
    orp, ocol = bo.transpose_csr(n, rowptr, col)
    assert np.array_equal(trp, orp) and np.array_equal(tcol, ocol)
    brp, bcol = _transpose_on_gpu(cnr, n, trp, tcol)
    assert np.array_equal(brp, rowptr) and np.array_equal(bcol, col)
    rng = np.random.default_rng(3)
    for nn, m in ((1, 0), (5, 0), (300, 2000), (70000, 300000), (300000, 1200000)):
        pairs = np.unique(np.stack([rng.integers(0, nn, m), rng.integers(0, nn, m)], axis=1), axis=0) if m else np.zeros((0, 2), np.int64)
        rp = np.zeros(nn + 1, dtype=np.int64)
        np.cumsum(np.bincount(pairs[:, 0], minlength=nn), out=rp[1:])
        c = pairs[:, 1].astype(np.int64)
        trp, tcol = _transpose_on_gpu(cnr, nn, rp, c)
        orp, ocol = bo.transpose_csr(nn, rp, c)
        assert np.array_equal(trp, orp) and np.array_equal(tcol, ocol), (nn, m)
    star_rp = np.arange(0, 50001, dtype=np.int64)
    star_rp[0] = 0
    star_col = np.full(50000, 7, dtype=np.int64)
    trp, tcol = _transpose_on_gpu(cnr, 50000, star_rp, star_col)
    assert trp[7] == 0 and trp[8] == 50000 and np.array_equal(tcol, np.arange(50000))
    with pytest.raises(wg.IllegalArgumentException):
        _transpose_on_gpu(cnr, 10, np.array([0, 1] + [1] * 9, dtype=np.int64), np.array([10], dtype=np.int64))


def test_symmetrize_equals_oracle_and_the_union_with_the_transpose(cnr, cnr_golden):
    """wgbv_symmetrize_csr against the CPU restatement of Transform.symmetrizeOffline on the golden fixture;
    the reference's own definition symmetrize(g) == union(g, transpose(g)) (Transform.java:573-574) through the separate
    entry points; the result is its own transpose; a graph with loops, empty rows and duplicate arcs between g and g^T."""
    torch = pytest.importorskip("torch")
    deg, rowptr, col = cnr_golden
    n = cnr.numNodes()

    def sym(nn, rp, c):
        d_rp = torch.from_numpy(np.ascontiguousarray(rp, dtype=np.int64)).cuda()
        d_c = torch.from_numpy(np.ascontiguousarray(c, dtype=np.int64)).cuda() if c.size else torch.zeros(1, dtype=torch.int64, device="cuda")
        d_srp = torch.empty(nn + 1, dtype=torch.int64, device="cuda")
        arcs = cnr.symmetrizeCSRDevice(nn, d_rp.data_ptr(), d_c.data_ptr(), d_srp.data_ptr())
        d_sc = torch.empty(max(arcs, 1), dtype=torch.int64, device="cuda")
        assert cnr.symmetrizeCSRDevice(nn, d_rp.data_ptr(), d_c.data_ptr(), d_srp.data_ptr(), d_sc.data_ptr(), arcs) == arcs
        return d_srp.cpu().numpy(), d_sc.cpu().numpy()[:arcs]

    srp, scol = sym(n, rowptr, col)
    orp, ocol = bo.symmetrize_csr(n, rowptr, col)
    assert np.array_equal(srp, orp) and np.array_equal(scol, ocol)
    # union(g, transpose(g)) through the two separate entry points
    trp, tcol = _transpose_on_gpu(cnr, n, rowptr, col)
    t = [torch.from_numpy(np.ascontiguousarray(a, dtype=np.int64)).cuda() for a in (rowptr, col, trp, tcol)]
    d_urp = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    d_uc = torch.empty(scol.size, dtype=torch.int64, device="cuda")
    assert cnr.unionCSRDevice(n, t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr(), t[3].data_ptr(), d_urp.data_ptr(), d_uc.data_ptr(), scol.size) == scol.size
    assert np.array_equal(d_urp.cpu().numpy(), srp) and np.array_equal(d_uc.cpu().numpy(), scol)
    with pytest.raises(wg.IllegalArgumentException):
        cnr.unionCSRDevice(n, t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr(), t[3].data_ptr(), d_urp.data_ptr(), d_uc.data_ptr(), scol.size - 1)
    # symmetric: equal to its own transpose
    s2rp, s2col = _transpose_on_gpu(cnr, n, srp, scol)
    assert np.array_equal(s2rp, srp) and np.array_equal(s2col, scol)
    rng = np.random.default_rng(9)
    for nn, m in ((1, 0), (4, 0), (50, 400), (20000, 90000)):
        pairs = np.unique(np.stack([rng.integers(0, nn, m), rng.integers(0, nn, m)], axis=1), axis=0) if m else np.zeros((0, 2), np.int64)
        rp = np.zeros(nn + 1, dtype=np.int64)
        np.cumsum(np.bincount(pairs[:, 0], minlength=nn), out=rp[1:])
        c = pairs[:, 1].astype(np.int64)
        srp, scol = sym(nn, rp, c)
        orp, ocol = bo.symmetrize_csr(nn, rp, c)
        assert np.array_equal(srp, orp) and np.array_equal(scol, ocol), (nn, m)


def test_offsets_decoded_on_the_device(monkeypatch, cnr_basename, cnr_golden):
    """the .offsets stream decoded and turned into the Elias-Fano list on the device (forced here for small
    graphs with WGBV_DEVICE_OFFSETS=1) must open the same graph as the host decoder: the reference's own .offsets file,
    graphs with giant records (sparse Elias-Fano blocks, codes longer than a machine word apart), zero-degree runs, a
    slice, and a truncated stream."""
    monkeypatch.setenv("WGBV_DEVICE_OFFSETS", "1")
    deg, rowptr, col = cnr_golden
    with wg.GpuBVGraph.load(cnr_basename) as g:
        rp, c = g.decodeToCSR()
        assert np.array_equal(rp, rowptr) and np.array_equal(c, col)
        ids = np.array([0, 1, 325556, 100000, 77, 200001], dtype=np.int64)
        rp2, c2 = g.successorsBatch(ids)
        assert np.array_equal(c2, np.concatenate([col[rowptr[x]:rowptr[x + 1]] for x in ids]))
    graphs = [bg.copy_model(50000, seed=31, mean_degree=9, nsplits=1),
              bg.erdos_renyi(3000, 0.01, seed=4),
              bg.big_graph(2000, giant_every=300, giant_degree=40000, nsplits=1)]
    rowptr2, col2 = _similar_rows_graph(3000, (256, 1000, 1536), 12, 40000, 40000)
    graphs.append(bg.encode_csr(rowptr2, col2, bg.make_params(window=7, maxref=3, minint=4)))
    empty_rp = np.zeros(5001, dtype=np.int64)
    empty_rp[2500:] = 3
    graphs.append(bg.encode_csr(empty_rp, np.array([1, 2, 4999], dtype=np.int64), bg.make_params()))
    for e in graphs:
        if e is None:
            continue
        odeg, ocol = oracle_of(e).decode_range()
        with gpu_of(e) as g:
            rp, c = g.decodeToCSR()
            assert np.array_equal(rp, rowptr_of(odeg)) and np.array_equal(c, ocol)
            n = odeg.size
            ids = np.array([0, n - 1, n // 2, n // 3], dtype=np.int64)
            assert np.array_equal(g.outdegrees(ids), odeg[ids])
    e = graphs[0]
    with pytest.raises(wg.IOException):
        wg.GpuBVGraph.load_mem(e.properties_text(), e.graph, e.offsets_stream[: e.offsets_stream.size // 2])


def test_randomized_parity_sweep():
    """Many small graphs with random shapes and random format parameters (window, maxrefcount, minintervallength,
    degrees up to a few thousand, how much is copied): bulk decode, a decode that starts in the middle (halo), batched
    random access and the transpose involution against the oracle."""
    rng = np.random.default_rng(20241)
    for trial in range(36):
        n = int(rng.integers(500, 40000))
        window = int(rng.choice([1, 2, 3, 7, 7, 7, 12, 16]))
        maxref = int(rng.choice([1, 2, 3, 3, 3, 5, 8, 2147483647]))
        minint = int(rng.choice([0, 2, 3, 4, 4, 8]))
        md = float(rng.choice([2, 5, 12, 20, 20, 45, 90]))
        e = bg.copy_model(n, seed=int(rng.integers(1, 1 << 30)), mean_degree=md, p_ref=float(rng.uniform(0.2, 0.95)),
                          p_keep=float(rng.uniform(0.3, 0.97)), p_run=float(rng.uniform(0.0, 0.6)), max_degree=int(rng.choice([50, 300, 3000])),
                          params=bg.make_params(window=window, maxref=maxref, minint=minint), nsplits=int(rng.choice([1, 3])))
        odeg, ocol = oracle_of(e).decode_range()
        orp = rowptr_of(odeg)
        tag = (trial, n, window, maxref, minint, md)
        with gpu_of(e) as g:
            rp, c = g.decodeToCSR()
            assert np.array_equal(rp, orp) and np.array_equal(c, ocol), tag
            a = int(rng.integers(0, n))
            b = int(rng.integers(a, n + 1))
            rp2, c2 = g.decodeToCSR(a, b)
            assert np.array_equal(rp2, orp[a:b + 1] - orp[a]) and np.array_equal(c2, ocol[orp[a]:orp[b]]), tag
            ids = rng.integers(0, n, 200).astype(np.int64)
            rp3, c3 = g.successorsBatch(ids)
            assert np.array_equal(c3, np.concatenate([ocol[orp[x]:orp[x + 1]] for x in ids])), tag
            if trial % 6 == 0:
                trp, tcol = _transpose_on_gpu(g, n, orp, ocol)
                brp, bcol = _transpose_on_gpu(g, n, trp, tcol)
                assert np.array_equal(brp, orp) and np.array_equal(bcol, ocol), tag


def test_arc_labels_equal_oracle(cnr, cnr_golden):
    """.labels / .labeloffsets of a BitStreamArcLabelledImmutableGraph decoded on the device, against the
    CPU restatement: fixed-width labels of several widths (0, 1, 10, 31, 32 bits: the last one exercises the int sign),
    gamma-coded labels, ranges that start in the middle, and a label stream that does not match the graph."""
    deg, rowptr, col = cnr_golden
    n = deg.size
    rng = np.random.default_rng(11)
    for kind, width, hi in ((wg.LABEL_FIXED_INT, 10, 1 << 10), (wg.LABEL_FIXED_INT, 1, 2), (wg.LABEL_FIXED_INT, 0, 1),
                            (wg.LABEL_FIXED_INT, 31, 1 << 31), (wg.LABEL_FIXED_INT, 32, 1 << 32), (wg.LABEL_GAMMA_INT, 0, 1 << 18)):
        values = rng.integers(0, hi, col.size).astype(np.int64)
        lab, off = bg.write_labels(kind, width, rowptr, values)
        expect = bo.decode_labels(kind, width, lab, off, deg)
        assert np.array_equal(expect.astype(np.uint32), values.astype(np.uint32))
        with wg.GpuArcLabels(kind, width, lab, off, n) as L:
            got = L.decode(cnr, rowptr)
            assert np.array_equal(got, expect), (kind, width)
            a, b = 123456, 200001
            rp, _ = cnr.decodeToCSR(a, b)
            assert np.array_equal(L.decode(cnr, rp, a, b), expect[rowptr[a]:rowptr[b]])
            assert L.decode(cnr, np.zeros(1, dtype=np.int64), 5, 5).size == 0
            wrong = rowptr.copy()
            wrong[1:] += 1  # node 0 claims one more arc than its labels hold
            if not (kind == wg.LABEL_FIXED_INT and width == 0):  # (labels of no bits end where they should, whatever their number)
                with pytest.raises(wg.IllegalStateException):
                    L.decode(cnr, wrong[:1001], 0, 1000)
            with pytest.raises(wg.IllegalArgumentException):
                L.decode(cnr, rowptr[:3], n - 1, n + 1)


def test_list_labels_equal_oracle(cnr, cnr_golden):
    """SURVEY 8f-3, list-valued prototypes (FixedWidthIntListLabel.java:73-78, FixedWidthLongListLabel.java:78-84): per arc
    a gamma-coded length and that many fixed-width elements; the device returns the list starts and the elements.  Int
    lists of 0, 7 and 32 bits (sign), long lists of 33, 63 and 64 bits, empty lists, sub-ranges, a mismatched stream."""
    deg, rowptr, col = cnr_golden
    n = 60000
    rp = rowptr[: n + 1]
    arcs = int(rp[-1])
    rng = np.random.default_rng(5)
    for kind, width in ((wg.LABEL_FIXED_INT_LIST, 7), (wg.LABEL_FIXED_INT_LIST, 0), (wg.LABEL_FIXED_INT_LIST, 32),
                        (wg.LABEL_FIXED_LONG_LIST, 33), (wg.LABEL_FIXED_LONG_LIST, 63), (wg.LABEL_FIXED_LONG_LIST, 64)):
        lens = rng.integers(0, 6, arcs)
        lens[rng.integers(0, arcs, 20)] = 300  # a few long lists
        lp = np.zeros(arcs + 1, dtype=np.int64)
        np.cumsum(lens, out=lp[1:])
        raw = rng.integers(0, 1 << 63, int(lp[-1]), dtype=np.uint64) * 2 + rng.integers(0, 2, int(lp[-1]), dtype=np.uint64)
        values = (raw >> np.uint64(64 - width) if width else raw * np.uint64(0)).view(np.int64)
        lab, off = bg.write_label_lists(width, rp, lp, values)
        elp, evals = bo.decode_label_lists(kind, width, lab, off, deg[:n])
        assert np.array_equal(elp, lp)
        want = values if kind == wg.LABEL_FIXED_LONG_LIST else values.astype(np.uint32).view(np.int32).astype(np.int64)
        assert np.array_equal(evals, want)
        with wg.GpuArcLabels(kind, width, lab, off, n) as L:
            glp, gv = L.decodeLists(cnr, rp)
            assert np.array_equal(glp, elp) and np.array_equal(gv, evals), (kind, width)
            a, b = 12345, 40001
            r2, _ = cnr.decodeToCSR(a, b)
            glp, gv = L.decodeLists(cnr, r2, a, b)
            assert np.array_equal(glp, elp[rp[a]:rp[b] + 1] - elp[rp[a]]) and np.array_equal(gv, evals[elp[rp[a]]:elp[rp[b]]])
            glp, gv = L.decodeLists(cnr, np.zeros(1, dtype=np.int64), 5, 5)
            assert glp.tolist() == [0] and gv.size == 0
            with pytest.raises(wg.UnsupportedOperationException):
                L.decode(cnr, rp)
            wrong = rp.copy()
            wrong[1:] += 1
            with pytest.raises(wg.IllegalStateException):
                L.decodeLists(cnr, wrong[:1001], 0, 1000)


def test_device_compressor_reproduces_the_reference_fixture(cnr_basename, cnr_golden):
    """the device compressor (wgbv_encode_csr) given the golden successor lists and the fixture's parameters
    (w=7, maxRef=3, minInt=3, zeta_3; one run, as the fixture was stored) writes the reference's own cnr-2000.graph byte
    for byte, and record lengths whose gamma-coded differences are cnr-2000.offsets."""
    deg, rowptr, col = cnr_golden
    graph, offsets, props = wg.store_csr(rowptr, col, window_size=7, max_ref_count=3, min_interval_length=3, zeta_k=3)
    want = np.fromfile(cnr_basename + ".graph", dtype=np.uint8)
    assert graph.size == want.size and np.array_equal(graph, want)
    want_off = np.fromfile(cnr_basename + ".offsets", dtype=np.uint8)
    assert np.array_equal(offsets, want_off)


@pytest.mark.parametrize("window,maxref,minint,split", [(7, 3, 4, 0), (7, 3, 4, 7000), (1, 1, 0, 0), (0, 3, 2, 0), (16, 8, 1, 5000), (3, 0, 3, 0)])
def test_device_compressor_equals_the_generator_and_round_trips(window, maxref, minint, split):
    """The device compressor against the CPU restatement of the reference's compressor (datagen/bvgen.c, bit-exact on the
    fixture) for other parameters and for compression splits; what it writes decodes back to the same lists."""
    e0 = bg.copy_model(40000, seed=900 + window, mean_degree=14, nsplits=1)
    odeg, ocol = oracle_of(e0).decode_range()
    rp = rowptr_of(odeg)
    e = bg.encode_csr(rp, ocol, bg.make_params(window=window, maxref=maxref, minint=minint), nsplits=-split if split else 1)
    graph, offsets, props = wg.store_csr(rp, ocol, window_size=window, max_ref_count=maxref, min_interval_length=minint, split_nodes=split)
    assert np.array_equal(graph, e.graph)
    assert np.array_equal(offsets, e.offsets_stream)
    with wg.GpuBVGraph.load_mem(props, graph, offsets) as g:
        rp2, c2 = g.decodeToCSR()
        assert np.array_equal(rp2, rp) and np.array_equal(c2, ocol)
