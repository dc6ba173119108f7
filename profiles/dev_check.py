"""Development check of the bulk decode (run on the GPU box): decodes a few graphs through the C ABI, compares with the
oracle and, on a mismatch, says where (row, tile, position, what was expected) instead of just failing.

  python profiles/dev_check.py [--big]
"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "datagen")):
    sys.path.insert(0, p)

import numpy as np  # noqa: E402

import bvgen as bg  # noqa: E402
import bvoracle  # noqa: E402
import webgraph_big_b200 as wg  # noqa: E402


def oracle_csr(e):
    p = bvoracle.params_from_properties(bvoracle.parse_properties(e.properties_text()))
    deg, col = bvoracle.OracleGraph(p, e.graph, e.offsets_stream.tobytes()).decode_range()
    rp = np.zeros(deg.size + 1, dtype=np.int64)
    np.cumsum(deg, out=rp[1:])
    return rp, col


def report(name, rp, col, orp, ocol, tile_rows):
    if np.array_equal(rp, orp) and np.array_equal(col, ocol):
        return True
    print(f"[{name}] MISMATCH")
    if not np.array_equal(rp, orp):
        bad = np.nonzero(rp != orp)[0]
        print(f"  rowptr differs at {bad.size} entries, first {bad[:5]}, tile {bad[0] // tile_rows}; got {rp[bad[:5]]} want {orp[bad[:5]]}")
    m = min(col.size, ocol.size)
    bad = np.nonzero(col[:m] != ocol[:m])[0]
    print(f"  col sizes {col.size} / {ocol.size}; {bad.size} differing positions")
    if bad.size:
        rows = np.searchsorted(orp, bad, side="right") - 1
        urows = np.unique(rows)
        print(f"  {urows.size} rows affected; first rows {urows[:10]} (tiles {np.unique(urows[:50] // tile_rows)[:10]})")
        for r in urows[:4]:
            a, b = orp[r], orp[r + 1]
            print(f"   row {r} (tile {r // tile_rows}, index in tile {r % tile_rows}) degree {b - a}\n     want {ocol[a:b][:24]}\n     got  {col[a:min(b, col.size)][:24]}")
    return False


def check(name, e, ranges=((None, None),), **kw):
    orp, ocol = oracle_csr(e)
    ok = True
    with wg.GpuBVGraph.load_mem(e.properties_text(), e.graph, e.offsets_stream, **kw) as g:
        halo = max(e.params.window, 0) * max(e.params.maxref, 0)
        tile_rows = 256 - halo if 0 <= halo <= 64 else 256
        for lo, hi in ranges:
            lo = 0 if lo is None else lo
            hi = e.n if hi is None else hi
            t0 = time.time()
            rp, col = g.decodeToCSR(lo, hi)
            dt = time.time() - t0
            st = g.fusedStats()
            good = report(f"{name} [{lo},{hi})", rp, col, orp[lo:hi + 1] - orp[lo], ocol[orp[lo]:orp[hi]], tile_rows)
            print(f"[{name}] [{lo},{hi}) {'ok' if good else 'FAILED'}  {hi - lo} nodes {col.size} arcs  {dt * 1e3:.1f} ms  {st}", flush=True)
            ok = ok and good
    return ok


def main():
    ok = True
    e = bg.copy_model(3000, seed=1, mean_degree=8, nsplits=1)
    ok &= check("tiny", e, ranges=((None, None), (0, 1), (100, 101), (235, 236), (234, 471), (1000, 3000)))
    e = bg.copy_model(100000, seed=2, mean_degree=20, nsplits=4)
    ok &= check("copy100k", e, ranges=((None, None), (12345, 77777)))
    for w, m, i in ((0, 3, 4), (1, 1, 0), (3, 0, 2), (8, 8, 1), (7, 3, 3)):
        e = bg.copy_model(40000, seed=10 + w, mean_degree=15, params=bg.make_params(window=w, maxref=m, minint=i), nsplits=2)
        ok &= check(f"w{w}m{m}i{i}", e)
    e = bg.copy_model(50000, seed=5, mean_degree=60, nsplits=2)
    ok &= check("dense60", e)
    e = bg.copy_model(20000, seed=6, mean_degree=300, max_degree=20000, nsplits=2)
    ok &= check("dense300", e)
    e = bg.big_graph(70000, giant_every=30000, giant_degree=50000, giant_stride=1)
    ok &= check("giants", e)
    e = bg.big_graph(70000, giant_every=20000, giant_degree=60000, giant_stride=4)
    ok &= check("giants4", e)
    for res in (bg.GAMMA, bg.DELTA, bg.NIBBLE, bg.GOLOMB):
        e = bg.copy_model(30000, seed=31 + res, mean_degree=40, params=bg.make_params(residual=res, zetak=5 if res == bg.GOLOMB else 3), nsplits=2)
        ok &= check(f"res{res}", e)
    e = bg.copy_model(30000, seed=77, mean_degree=40, p_run=0.9, params=bg.make_params(minint=2), nsplits=2)
    ok &= check("intervals", e)
    e = bg.erdos_renyi(50000, p=2e-4, seed=1) if hasattr(bg, "erdos_renyi") else None
    if e is not None:
        ok &= check("er", e)
    if "--big" in sys.argv:
        e = bg.copy_model(5_000_000, seed=2, mean_degree=20, nsplits=32)
        ok &= check("copy5M", e)
    print("ALL OK" if ok else "SOME FAILED")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
