"""Random access (BASELINE config 4 at one fifth of the size): 2e6 random ids on a 2e7-node graph, for an ncu launch list."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "datagen")):
    sys.path.insert(0, p)
import torch
import bvgen as bg
import webgraph_big_b200 as wg
n = 20_000_000
e = bg.copy_model(n, seed=2, mean_degree=20, nsplits=-400000, threads=16)
dev = torch.device("cuda:0")
k = 2_000_000
gen = torch.Generator(device="cpu"); gen.manual_seed(3)
ids = torch.randint(0, n, (k,), generator=gen, dtype=torch.int64).to(dev)
with wg.GpuBVGraph.load_mem(e.properties_text(), e.graph, e.offsets_stream) as g:
    b_rp = torch.empty(k + 1, dtype=torch.int64, device=dev)
    arcs = g.successorsBatchDevice(ids.data_ptr(), k, b_rp.data_ptr())
    b_col = torch.empty(max(arcs, 1), dtype=torch.int64, device=dev)
    for _ in range(2):
        g.successorsBatchDevice(ids.data_ptr(), k, b_rp.data_ptr(), b_col.data_ptr(), arcs)
    g.sync()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    ev[0].record()
    for _ in range(3):
        g.successorsBatchDevice(ids.data_ptr(), k, b_rp.data_ptr(), b_col.data_ptr(), arcs)
    ev[1].record(); g.sync(); torch.cuda.synchronize()
    print("c4 probe: %d queries, %d arcs, %.3f ms" % (k, arcs, ev[0].elapsed_time(ev[1]) / 3), flush=True)
