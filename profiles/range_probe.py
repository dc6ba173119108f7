"""One range decode of 250k nodes in the middle of a 5M-node graph (for an ncu launch list of the per-call fixed cost)."""
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
    for _ in range(3):
        g.decodeToCSRDevice(2_000_000, 2_250_000, d_rp.data_ptr(), d_col.data_ptr(), e.arcs, sync=True)
