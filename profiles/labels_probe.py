"""ms per decode of gamma-coded labels on a 2e7-node graph (values below 2^18), and the check against the generator's values."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "datagen")):
    sys.path.insert(0, p)
import numpy as np, torch
import bvgen as bg
import webgraph_big_b200 as wg
n = 20_000_000
e = bg.copy_model(n, seed=2, mean_degree=20, nsplits=-400000, threads=16)
dev = torch.device("cuda:0")
with wg.GpuBVGraph.load_mem(e.properties_text(), e.graph, e.offsets_stream) as g:
    d_rp = torch.empty(n + 1, dtype=torch.int64, device=dev)
    d_col = torch.empty(e.arcs, dtype=torch.int64, device=dev)
    g.decodeToCSRDevice(0, n, d_rp.data_ptr(), d_col.data_ptr(), e.arcs)
    rp = d_rp.cpu().numpy()
    rng = np.random.default_rng(1)
    vals = rng.integers(0, 1 << 18, e.arcs).astype(np.int64)
    lab, off = bg.write_labels(bg.LABEL_GAMMA_INT, 0, rp, vals)
    with wg.GpuArcLabels(wg.LABEL_GAMMA_INT, 0, lab, off, n) as L:
        d_out = torch.empty(e.arcs, dtype=torch.int64, device=dev)
        for _ in range(2):
            L.decodeDevice(g, 0, n, d_rp.data_ptr(), d_out.data_ptr(), e.arcs)
        g.sync()
        import time
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(5):
            L.decodeDevice(g, 0, n, d_rp.data_ptr(), d_out.data_ptr(), e.arcs)
        g.sync()
        ms = (time.perf_counter() - t0) / 5 * 1e3  # (the library launches on its own stream: wall clock around its sync)
        ok = bool(np.array_equal(d_out.cpu().numpy(), vals))
        print("%s gamma labels: %d labels, %.3f ms, equal=%s" % (os.environ.get("WGBV_LIBRARY", "default").split("/")[-1], e.arcs, ms, ok), flush=True)
