"""Does keeping a chunk's rows in L2 between the extras kernel and the copy levels pay?  Decodes a 20M-node graph once
as a whole and once as consecutive ranges of C nodes (each into its slice of the same arrays), CUDA-event timed."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "datagen")):
    sys.path.insert(0, p)
import numpy as np, torch
import bvgen as bg

import webgraph_big_b200 as wg

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20_000_000
e = bg.copy_model(n, seed=1, mean_degree=20, nsplits=-400000, threads=16)
dev = torch.device("cuda:0")
with wg.GpuBVGraph.load_mem(e.properties_text(), e.graph, e.offsets_stream) as g:
    arcs = e.arcs
    d_rp = torch.empty(n + 1 + 64, dtype=torch.int64, device=dev)
    d_col = torch.empty(arcs + 64, dtype=torch.int64, device=dev)
    def full():
        g.decodeToCSRDevice(0, n, d_rp.data_ptr(), d_col.data_ptr(), arcs, sync=False)
    for _ in range(3): full()
    g.sync()
    rp = d_rp[: n + 1].cpu().numpy()
    ref = d_col[:arcs].clone()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
    torch.cuda.synchronize(); ev[0].record()
    for _ in range(5): full()
    ev[1].record(); g.sync(); torch.cuda.synchronize()
    print("whole: %.3f ms" % (ev[0].elapsed_time(ev[1]) / 5), flush=True)
    for C in (250_000, 500_000, 1_000_000, 2_000_000, 5_000_000):
        bounds = list(range(0, n, C)) + [n]
        d_rp2 = torch.empty(C + 1 + 64, dtype=torch.int64, device=dev)
        def chunks():
            for a, b in zip(bounds[:-1], bounds[1:]):
                g.decodeToCSRDevice(a, b, d_rp2.data_ptr(), d_col.data_ptr() + 8 * int(rp[a]), int(rp[b] - rp[a]), sync=False)
        d_col.zero_()
        chunks(); g.sync(); torch.cuda.synchronize()
        ok = bool(torch.equal(d_col[:arcs], ref))
        torch.cuda.synchronize(); t0 = time.perf_counter(); ev[0].record()
        for _ in range(3): chunks()
        ev[1].record(); g.sync(); torch.cuda.synchronize()
        print("chunks of %8d: %.3f ms (wall %.3f) equal=%s" % (C, ev[0].elapsed_time(ev[1]) / 3, (time.perf_counter() - t0) / 3 * 1e3, ok), flush=True)
