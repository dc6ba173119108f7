"""ms per range decode of 250k / 1M nodes in the middle of a 5M-node graph, and the whole graph (library = WGBV_LIBRARY)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "datagen")):
    sys.path.insert(0, p)
import torch
import bvgen as bg
import webgraph_big_b200 as wg
n = 5_000_000
e = bg.copy_model(n, seed=1, mean_degree=20, nsplits=-400000, threads=16)
dev = torch.device("cuda:0")
with wg.GpuBVGraph.load_mem(e.properties_text(), e.graph, e.offsets_stream) as g:
    d_rp = torch.empty(n + 65, dtype=torch.int64, device=dev)
    d_col = torch.empty(e.arcs + 64, dtype=torch.int64, device=dev)
    out = []
    for a, b in ((2_000_000, 2_250_000), (2_000_000, 3_000_000), (0, n)):
        for _ in range(3):
            g.decodeToCSRDevice(a, b, d_rp.data_ptr(), d_col.data_ptr(), e.arcs, sync=True)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        torch.cuda.synchronize(); ev[0].record()
        for _ in range(10):
            g.decodeToCSRDevice(a, b, d_rp.data_ptr(), d_col.data_ptr(), e.arcs, sync=False)
        ev[1].record(); g.sync(); torch.cuda.synchronize()
        out.append("%d nodes: %.3f ms" % (b - a, ev[0].elapsed_time(ev[1]) / 10))
    print(os.environ.get("WGBV_LIBRARY", "default").split("/")[-1], " | ".join(out), flush=True)
