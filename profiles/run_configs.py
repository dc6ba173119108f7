"""Measures the BASELINE.json configurations other than the headline one (configs[1], which bench.py measures):
C1 (ER 1M nodes, CPU oracle full nodeIterator decode), C3 (copy-model graph recompressed with unbounded reference
chains, bulk decode on one B200) and C4 (batched random access, 10M random node ids on the C2 file).
Prints one JSON line per configuration.  usage: python profiles/run_configs.py [--nodes N] [--queries Q]"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "datagen")):
    sys.path.insert(0, p)
import numpy as np  # noqa: E402

import bvgen  # noqa: E402
import bvoracle  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nodes", type=int, default=100_000_000)
    ap.add_argument("--queries", type=int, default=10_000_000)
    ap.add_argument("--skip", default="")
    args = ap.parse_args()
    import torch

    import webgraph_big_b200 as wg
    threads = len(os.sched_getaffinity(0))
    torch.cuda.set_stream(torch.cuda.Stream())  # an explicit stream: the library launches on it, the events below are recorded on it

    if "c1" not in args.skip:
        e = bvgen.erdos_renyi(1_000_000, 1.6e-5, seed=1, params=bvgen.make_params(7, 3, 4, 3))
        p = bvoracle.params_from_properties(bvoracle.parse_properties(e.properties_text()))
        og = bvoracle.OracleGraph(p, e.graph, offsets=e.offsets)
        t0 = time.time()
        deg, _ = og.decode_range(0, e.n, threads=1, want_col=False)
        dt = time.time() - t0
        with wg.GpuBVGraph.load_mem(e.properties_text(), e.graph, e.offsets_stream) as g:
            d_rp = torch.empty(e.n + 1, dtype=torch.int64, device="cuda")
            d_col = torch.empty(e.arcs, dtype=torch.int64, device="cuda")
            g.decodeToCSRDevice(0, e.n, d_rp.data_ptr(), d_col.data_ptr(), e.arcs)
            ok = g.csrChecksumDevice(0, e.n, d_rp.data_ptr(), d_col.data_ptr()) == e.checksum
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            for _ in range(10):
                g.decodeToCSRDevice(0, e.n, d_rp.data_ptr(), d_col.data_ptr(), e.arcs)
            ev1.record()
            torch.cuda.synchronize()
            gms = ev0.elapsed_time(ev1) / 10
        print(json.dumps({"config": "C1 ER 1M nodes w=7 maxRef=3 minInt=4 zeta3", "arcs": int(e.arcs), "bits_per_arc": e.graph_bits / e.arcs,
                          "cpu_oracle_sequential_arcs_per_s": e.arcs / dt, "gpu_ms": gms, "gpu_arcs_per_s": e.arcs / (gms * 1e-3), "verified": bool(ok)}))

    n = args.nodes
    if "c3" not in args.skip:
        t0 = time.time()
        e = bvgen.copy_model(n, seed=2, params=bvgen.make_params(maxref=2147483647), nsplits=256)
        gen = time.time() - t0
        with wg.GpuBVGraph.load_mem(e.properties_text(), e.graph, e.offsets_stream, cuda_stream=torch.cuda.current_stream().cuda_stream) as g:
            d_rp = torch.empty(n + 1, dtype=torch.int64, device="cuda")
            d_col = torch.empty(e.arcs, dtype=torch.int64, device="cuda")
            for _ in range(2):
                g.decodeToCSRDevice(0, n, d_rp.data_ptr(), d_col.data_ptr(), e.arcs)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record()
            for _ in range(5):
                g.decodeToCSRDevice(0, n, d_rp.data_ptr(), d_col.data_ptr(), e.arcs)
            ev1.record()
            torch.cuda.synchronize()
            ms = ev0.elapsed_time(ev1) / 5
            ok = g.csrChecksumDevice(0, n, d_rp.data_ptr(), d_col.data_ptr()) == e.checksum
            print(json.dumps({"config": f"C3 copy-model {n} nodes, maxrefcount=2147483647", "arcs": int(e.arcs), "bits_per_arc": e.graph_bits / e.arcs,
                              "ms_per_decode": ms, "arcs_per_s": e.arcs / (ms * 1e-3), "stage_ms": g.lastStageMs(), "verified": bool(ok), "generation_s": gen}))
        del d_rp, d_col

    if "c4" not in args.skip:
        e = bvgen.copy_model(n, seed=2, nsplits=256)
        p = bvoracle.params_from_properties(bvoracle.parse_properties(e.properties_text()))
        og = bvoracle.OracleGraph(p, e.graph, offsets=e.offsets)
        with wg.GpuBVGraph.load_mem(e.properties_text(), e.graph, e.offsets_stream, cuda_stream=torch.cuda.current_stream().cuda_stream) as g:
            k = args.queries
            gen = torch.Generator(device="cuda").manual_seed(3)
            ids = torch.randint(0, n, (k,), generator=gen, device="cuda", dtype=torch.int64)
            d_rp = torch.empty(k + 1, dtype=torch.int64, device="cuda")
            arcs = g.successorsBatchDevice(ids.data_ptr(), k, d_rp.data_ptr())
            d_col = torch.empty(arcs, dtype=torch.int64, device="cuda")
            for _ in range(2):
                g.successorsBatchDevice(ids.data_ptr(), k, d_rp.data_ptr(), d_col.data_ptr(), arcs)
            t0 = time.perf_counter()
            reps = 5
            for _ in range(reps):
                g.successorsBatchDevice(ids.data_ptr(), k, d_rp.data_ptr(), d_col.data_ptr(), arcs)
            torch.cuda.synchronize()
            dt = (time.perf_counter() - t0) / reps
            # parity on a sample of the queries against the CPU oracle (recursive random access)
            h_ids = ids[:2000].cpu().numpy()
            h_rp = d_rp[:2001].cpu().numpy()
            h_col = d_col[: int(h_rp[-1])].cpu().numpy()
            ok = all(np.array_equal(og.successors(int(x)), h_col[h_rp[i]:h_rp[i + 1]]) for i, x in enumerate(h_ids))
            # CPU: the oracle's random access on one core, bounded sample
            t0 = time.time()
            m = 0
            for x in h_ids:
                m += og.successors(int(x)).size
            cdt = time.time() - t0
            print(json.dumps({"config": f"C4 {k} random successors(x) on copy-model {n} nodes", "arcs_out": int(arcs), "ms_per_batch": dt * 1e3,
                              "nodes_per_s": k / dt, "arcs_per_s": arcs / dt, "verified_sample": bool(ok),
                              "cpu_oracle_1core_nodes_per_s_python_loop": len(h_ids) / cdt}))


if __name__ == "__main__":
    main()
