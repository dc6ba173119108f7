#!/usr/bin/env python
"""bench.py -- decoded arcs/s of the BVGraph -> CSR hot path (BASELINE.json metric).

A "step" is one complete bulk decode of the rank's graph: compressed bit stream + Elias-Fano offsets resident in
HBM -> int64 CSR row pointers and successors resident in HBM.  The workload at N=1 is BASELINE.json configs[1]:
a synthetic copy-model web-like graph, 100M nodes / ~2B arcs, default BVGraph parameters (w=7, maxRef=3,
minInt=4, zeta_3), produced by this repository's generator + encoder (datagen/).  At N>1 every rank holds one
a shard of ONE N x 100M-node graph (global node ids).  Shard boundaries lie in the middle of compression splits, so the
first rows of a shard reference rows of the previous one: every rank opens its slice with the window x maxRef halo rows
before its first node and decodes them redundantly (no cross-GPU traffic), then the ranks all-gather their arc counts
over NCCL to form global CSR offsets: weak scaling.  --config c5 is BASELINE.json configs[4] (3e9 nodes over all ranks,
mean degree 3.3, giant rows); at N=1 the line also carries configs[0], [2] and [3] as sub-objects c1, c3, c4.

  python bench.py [--gpus N] [--steps K] [--warmup W]            our CUDA path
  python bench.py --impl reference [...]                          the reference's CPU algorithm (oracle port)

One JSON line is printed by rank 0.  See DESIGN.md ("Measurement") for how every field is obtained.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "datagen")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

METRIC = "decoded arcs/s (BVGraph->CSR)"
UNIT = "arcs/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--nodes", type=int, default=100_000_000, help="nodes per GPU (configs[1]: 100M)")
    ap.add_argument("--mean-degree", type=float, default=20.0)
    ap.add_argument("--maxref", type=int, default=3)
    ap.add_argument("--e2e-steps", type=int, default=2)
    ap.add_argument("--cpu-sample-nodes", type=int, default=0, help="0: sized from the core count")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-labels", action="store_true", help="skip the arc-label decode (SURVEY 8f-3, N=1 only)")
    ap.add_argument("--no-transpose", action="store_true", help="skip the transpose of the decoded CSR (SURVEY 8f-1, N=1 only)")
    ap.add_argument("--e2e-trace", action="store_true", help="print the arrival time of every chunk of the last e2e step to stderr")
    ap.add_argument("--seed", type=int, default=2)
    ap.add_argument("--max-degree", type=int, default=10000)
    ap.add_argument("--config", default="c2", choices=["c2", "c5"], help="c2: configs[1] (default); c5: configs[4], 3e9 nodes over all ranks")
    ap.add_argument("--total-nodes", type=int, default=3_000_000_000, help="--config c5: nodes of the whole graph")
    ap.add_argument("--giant-every", type=int, default=50_000_000, help="--config c5: one giant row every that many nodes")
    ap.add_argument("--giant-degree", type=int, default=2_000_000)
    ap.add_argument("--no-configs", action="store_true", help="skip the c1 / c3 / c4 sub-objects (N=1) and the c5 sub-object (N=8)")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def workload_name(args, world):
    if args.config == "c5":
        return (f"copy-model web-like graph with giant rows, {args.total_nodes / 1e9:g}e9 nodes over {world} GPU(s), mean degree {args.mean_degree:g}, "
                f"one row of {args.giant_degree} successors every {args.giant_every} nodes, BVGraph w=7 maxRef={args.maxref} minInt=4 zeta3, "
                "bulk decode to int64 CSR")
    return (f"copy-model web-like graph, {args.nodes // 1_000_000}M nodes/GPU x {world} GPU(s), mean degree {args.mean_degree:g}, "
            f"BVGraph w=7 maxRef={args.maxref} minInt=4 zeta3, bulk decode to int64 CSR")


def split_nodes(nodes_per_rank):
    """Nodes per compression split (even: shard boundaries lie in the middle of a split)."""
    return max(2, (nodes_per_rank // 250) & ~1)


def shard_bounds(args, world):
    """First node of every shard of the world x nodes graph: [0, n + P/2, 2n + P/2, ..., world * n]: boundaries in the
    middle of compression splits of P nodes, so that rows right after a boundary reference rows before it."""
    n = args.nodes
    P = split_nodes(n)
    return [0] + [(r * n) // P * P + P // 2 for r in range(1, world)] + [world * n]


class Shard:
    """What one rank holds: the compression splits of the global graph that cover its shard (generated and BV-encoded
    here, identical to what any other rank would generate for the same nodes), cut to a slice that starts halo rows
    before the shard."""

    def __init__(self, args, rank, world, threads, maxref=None):
        import bvgen
        t0 = time.time()
        P = split_nodes(args.nodes)
        b = shard_bounds(args, world)
        self.first, self.last = b[rank], b[rank + 1]         # the shard: nodes [first, last)
        self.part_first = (self.first // P) * P               # generation starts at the split that holds `first`
        self.n = self.last - self.first
        kw = {}
        if args.config == "c5":
            kw = dict(giant_every=args.giant_every, giant_degree=args.giant_degree, giant_stride=4)
        self.seg = P // 2 if world > 1 else 0
        params = bvgen.make_params(maxref=args.maxref if maxref is None else maxref)
        e = bvgen.copy_model(self.last - self.part_first, seed=args.seed, mean_degree=args.mean_degree, params=params, max_degree=args.max_degree,
                             nsplits=-P, threads=threads, first_node=self.part_first, global_n=world * args.nodes, seg_nodes=self.seg, **kw)
        self.e = e
        self.lead = self.first - self.part_first              # rows generated before the shard (half a split, or 0)
        # window x maxRef predecessors are all a row can reach (BVG:701, :2258)
        self.halo = min(self.lead, params.window * params.maxref) if params.maxref < 1000 else self.lead
        if self.lead:
            # the slice handed to the library: halo rows + the shard; its bits start at the record of the first halo row
            k0 = self.lead - self.halo
            self.checksum = int(np.sum(e.seg_checksums[1:], dtype=np.uint64))
            bit0 = int(e.offsets[k0])
            self.byte0 = bit0 >> 3
            self.graph = e.graph[self.byte0:]
            rel = (e.offsets[k0:] - np.uint64(self.byte0 * 8)).astype(np.uint64)
            self.offsets_stream = bvgen.write_codes(bvgen.GAMMA, np.diff(rel, prepend=np.uint64(0)))[0]
            self.slice_first = self.first - self.halo
        else:
            self.checksum = e.checksum
            self.graph = e.graph
            self.offsets_stream = e.offsets_stream
            self.slice_first = self.first
        self.gen_s = time.time() - t0

    def properties_text(self, world, nodes_total):
        return self.e.properties_text().replace(f"nodes={self.e.n}", f"nodes={nodes_total}")


class ClockSampler:
    """Samples SM clocks and throttle reasons with nvidia-smi while the timed region runs."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.out = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=self.out, stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        res = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return res
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.out.flush()
        self.out.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.out.read().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx.append(float(f[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        try:
            os.unlink(self.out.name)
        except OSError:
            pass
        if sm:
            res.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons), samples=len(sm))
        return res


def cpu_threads():
    return max(1, len(os.sched_getaffinity(0)))


def oracle_of(e):
    import bvoracle
    p = bvoracle.params_from_properties(bvoracle.parse_properties(e.properties_text()))
    # parts of a larger graph are decoded with local ids: the oracle only needs consistent offsets
    p.nodes = e.n
    return bvoracle.OracleGraph(p, e.graph, offsets=e.offsets)


def run_cpu(e, sample_nodes, threads, repeats=1):
    """The reference's CPU algorithm (oracle port of BVGraph.nodeIterator / splitNodeIterators) over the first
    `sample_nodes` nodes, all host threads; returns (arcs/s, arcs, seconds)."""
    og = oracle_of(e)
    best = None
    arcs = 0
    for _ in range(repeats):
        t0 = time.time()
        deg, _ = og.decode_range(0, sample_nodes, threads=threads, want_col=False)
        dt = time.time() - t0
        arcs = int(deg.sum())
        best = dt if best is None else min(best, dt)
    return arcs / best, arcs, best


def stats_of(e):
    """What the reference's compressor writes to .properties about a graph (BVGraph.java:2490-2567), from our encoder."""
    st = e.stats
    return {"copiedarcs": st.get("copiedarcs"), "intervalisedarcs": st.get("intervalisedarcs"), "residualarcs": st.get("residualarcs"),
            "rows_with_reference": st.get("rows_with_reference"), "avgref_chain": st.get("refchain_length_sum", 0) / max(st.get("rows_with_reference", 0), 1),
            "bitsforresiduals": st.get("bitsforresiduals"), "bitsforblocks": st.get("bitsforblocks"), "bitsforintervals": st.get("bitsforintervals"),
            "bitsforoutdegrees": st.get("bitsforoutdegrees"), "bitsforreferences": st.get("bitsforreferences")}


def main_reference(args, rank, world):
    """--impl reference: the reference's own algorithm on the host cores (no JVM on this image: the C restatement
    in oracle/, pinned on the reference's cnr-2000 fixture, run splitNodeIterators-style on all cores) over the WHOLE
    one-GPU workload, every step."""
    if rank != 0:
        return 0
    threads = cpu_threads()
    if args.config == "c5":
        args.nodes = min(args.total_nodes, 200_000_000)  # (the host decodes a 2e8-node part of the c5 graph per step)
        args.mean_degree = 3.3
    sh = Shard(args, 0, 1, threads)
    e = sh.e
    sample = args.cpu_sample_nodes or e.n
    for _ in range(max(args.warmup, 0)):
        run_cpu(e, min(sample, 500_000 * threads), threads)
    times, arcs = [], 0
    for _ in range(max(args.steps, 1)):
        rate, arcs, dt = run_cpu(e, sample, threads)
        times.append(dt)
    dt = float(np.mean(times))
    value = arcs / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": {"workload": workload_name(args, 1), "sample": f"all {sample} nodes ({arcs} arcs) per step" if sample == e.n else f"first {sample} nodes ({arcs} arcs) per step",
                   "warmup_sample_nodes": min(sample, 500_000 * threads)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{sample} of {e.n} nodes ({arcs} arcs), splitNodeIterators-style over {threads} threads, degrees + successor lists decoded"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "generation_s": sh.gen_s,
    }
    print(json.dumps(line))
    return 0


def time_decode(torch, g, first, last, d_rowptr, d_col, cap, steps, warmup, extra=None):
    """ms per bulk decode of [first, last) (CUDA events on the library's stream)."""
    for _ in range(warmup):
        g.decodeToCSRDevice(first, last, d_rowptr.data_ptr(), d_col.data_ptr(), cap, sync=False)
        if extra:
            extra()
    g.sync()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(steps):
        g.decodeToCSRDevice(first, last, d_rowptr.data_ptr(), d_col.data_ptr(), cap, sync=False)
        if extra:
            extra()
    ev1.record()
    g.sync()
    return ev0.elapsed_time(ev1) / steps


def pcie_probe(torch, dist, dev, world):
    """What bounds the end-to-end leg: pinned device->host copies of 1 GiB on every rank AT THE SAME TIME (and the same
    with a host->device copy beside each); GB/s per rank (slowest) and in total."""
    nbytes = 1 << 30
    d = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    h = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    d2 = torch.empty(nbytes // 4, dtype=torch.uint8, device=dev)
    h2 = torch.empty(nbytes // 4, dtype=torch.uint8, pin_memory=True)
    s2 = torch.cuda.Stream(device=dev)

    def run(with_h2d, reps=3):
        h.copy_(d, non_blocking=True)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            if with_h2d:
                with torch.cuda.stream(s2):
                    for _ in range(4):
                        d2.copy_(h2, non_blocking=True)
            h.copy_(d, non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        t = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return reps * nbytes / float(t.item()) / 1e9

    a, b = run(False), run(True)
    del d, h, d2, h2
    return {"d2h_gbs_per_rank_all_ranks_at_once": a, "d2h_gbs_total": a * world, "d2h_gbs_per_rank_with_h2d_beside": b,
            "what": "1 GiB pinned device-to-host copies on all ranks at the same time (slowest rank); the e2e leg moves 8 B per arc this way"}


def config_c1(torch, wg, dev, local_rank, stream):
    """BASELINE.json configs[0]: Erdos-Renyi 1M nodes / ~16M arcs, w=7 maxRef=3 minInt=4 zeta3, full nodeIterator decode on
    the CPU (the oracle's sequential scan, one thread) -- with the GPU decode of the same file beside it, compared in full."""
    import bvgen
    e = bvgen.erdos_renyi(1_000_000, p=1.6e-5, seed=1, nsplits=16)
    og = oracle_of(e)
    t0 = time.time()
    deg, col = og.decode_range(0, e.n, threads=1, want_col=True)
    cpu_s = time.time() - t0
    with wg.GpuBVGraph.load_mem(e.properties_text(), e.graph, e.offsets_stream, device=local_rank, cuda_stream=stream.cuda_stream) as g:
        d_rp = torch.empty(e.n + 1, dtype=torch.int64, device=dev)
        d_col = torch.empty(max(e.arcs, 1), dtype=torch.int64, device=dev)
        ms = time_decode(torch, g, 0, e.n, d_rp, d_col, e.arcs, 10, 3)
        rp = np.zeros(e.n + 1, dtype=np.int64)
        np.cumsum(deg, out=rp[1:])
        ok = bool(torch.equal(d_rp.cpu(), torch.from_numpy(rp))) and bool(torch.equal(d_col.cpu(), torch.from_numpy(col)))
        alg = g.algorithmicBytes(0, e.n, e.arcs)
    return {"workload": "Erdos-Renyi 1M nodes, p = 1.6e-5, BVGraph w=7 maxRef=3 minInt=4 zeta3", "nodes": e.n, "arcs": e.arcs, "bits_per_arc": e.graph_bits / max(e.arcs, 1),
            "cpu_sequential": {"value": e.arcs / cpu_s, "unit": UNIT, "seconds": cpu_s, "cores": 1, "kind": "port", "what": "oracle nodeIterator scan of the whole graph, successor lists materialised"},
            "gpu": {"value": e.arcs / (ms * 1e-3), "unit": UNIT, "ms": ms}, "verified": ok, "verified_how": "every row pointer and successor equal to the oracle's",
            "roofline": {"bound": "hbm", "achieved": alg / (ms * 1e-3) / 1e9, "unit": "GB/s", "frac": alg / (ms * 1e-3) / 1e9 / peaks()[0], "algorithmic_bytes": alg}}


def config_c3(torch, wg, args, dev, local_rank, stream, d_rowptr, d_col, n, arcs, checksum, threads):
    """configs[2]: the same graph recompressed with maxrefcount = 2147483647 (what -m -1 writes): unbounded reference chains,
    level-by-level copy resolution as deep as the longest chain.  The recompression is done by the device compressor
    (wgbv_encode_csr) from the CSR decoded above, so it is literally the same graph; what is decoded back must hash to the
    generator's checksum of that graph."""
    import bvgen
    P = split_nodes(args.nodes)
    t0 = time.time()
    out = np.zeros(arcs * 2 + n + 64, dtype=np.uint8)
    graph, bits, node_bits = wg.encode_csr_device(n, d_rowptr.data_ptr(), d_col.data_ptr(), arcs, max_ref_count=2147483647, split_nodes=P,
                                                  device=local_rank, out=out, want_node_bits=True)
    offsets_stream = bvgen.write_codes(bvgen.GAMMA, np.concatenate([np.zeros(1, dtype=np.uint64), node_bits[:n]]))[0]
    enc_s = time.time() - t0
    props = ("graphclass=it.unimi.dsi.big.webgraph.BVGraph\nversion=0\n"
             f"nodes={n}\narcs={arcs}\nwindowsize=7\nmaxrefcount=2147483647\nminintervallength=4\nzetak=3\ncompressionflags=\n")
    g = wg.GpuBVGraph.load_slice(props, 0, n, graph, offsets_stream, device=local_rank, cuda_stream=stream.cuda_stream)
    c_rowptr = torch.empty(n + 1, dtype=torch.int64, device=dev)
    c_col = torch.empty(max(arcs, 1), dtype=torch.int64, device=dev)
    l0 = g.kernelLaunches()
    ms = time_decode(torch, g, 0, n, c_rowptr, c_col, arcs, 3, 2)
    launches = (g.kernelLaunches() - l0) // 5
    ok = g.csrChecksumDevice(0, n, c_rowptr.data_ptr(), c_col.data_ptr()) == checksum and int(c_rowptr[n].item()) == arcs
    ok = ok and bool(torch.equal(c_rowptr, d_rowptr))
    alg = g.algorithmicBytes(0, n, arcs)
    g.close()
    del c_rowptr, c_col, out
    return {"workload": f"the headline graph ({n} nodes) recompressed on the device with maxrefcount=2147483647", "nodes": n, "arcs": arcs,
            "bits_per_arc": bits / max(arcs, 1), "ms": ms, "value": arcs / (ms * 1e-3), "unit": UNIT,
            "gpu_launches": int(launches), "verified": bool(ok),
            "verified_how": "checksum of every decoded row against the generator's checksum of the graph (taken before any encoding); row pointers equal to the headline decode's",
            "recompression_s": enc_s,
            "roofline": {"bound": "hbm", "achieved": alg / (ms * 1e-3) / 1e9, "unit": "GB/s", "frac": alg / (ms * 1e-3) / 1e9 / peaks()[0], "algorithmic_bytes": alg}}


def config_c4(torch, g, e, dev, d_rowptr, d_col, threads):
    """configs[3]: successors(x) for 1e7 random node ids (seed 3) through the Elias-Fano offsets, as one batch; checked
    against the bulk CSR decoded before; the CPU leg is the oracle's recursive successors(x) on all host threads
    (SpeedTest -r, test/SpeedTest.java:95-108) over the first 2e6 of the same ids."""
    n = e.n
    k = 10_000_000
    gen = torch.Generator(device="cpu")
    gen.manual_seed(3)
    ids = torch.randint(0, n, (k,), generator=gen, dtype=torch.int64)
    d_ids = ids.to(dev)
    b_rp = torch.empty(k + 1, dtype=torch.int64, device=dev)
    arcs = g.successorsBatchDevice(d_ids.data_ptr(), k, b_rp.data_ptr())
    b_col = torch.empty(max(arcs, 1), dtype=torch.int64, device=dev)
    g.successorsBatchDevice(d_ids.data_ptr(), k, b_rp.data_ptr(), b_col.data_ptr(), arcs)  # warm-up: sizes the scratch
    reps = 3
    l0 = g.kernelLaunches()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(reps):
        g.successorsBatchDevice(d_ids.data_ptr(), k, b_rp.data_ptr(), b_col.data_ptr(), arcs)
    ev1.record()
    g.sync()
    ms = ev0.elapsed_time(ev1) / reps
    launches = (g.kernelLaunches() - l0) // reps
    # check: row q of the batch must be row ids[q] of the CSR
    starts = d_rowptr[d_ids]
    lens = d_rowptr[d_ids + 1] - starts
    ok = bool(torch.equal(b_rp[1:] - b_rp[:-1], lens)) and int(b_rp[k].item()) == arcs
    if ok:
        for a in range(0, k, 1 << 20):
            bq = min(k, a + (1 << 20))
            ln = lens[a:bq]
            tot = int(ln.sum().item())
            src = torch.repeat_interleave(starts[a:bq] - (b_rp[a:bq] - b_rp[a]), ln) + torch.arange(tot, device=dev)
            lo = int(b_rp[a].item())
            ok = ok and bool(torch.equal(d_col[src], b_col[lo:lo + tot]))
            del src
    # CPU leg
    og = oracle_of(e)
    sample = ids[:2_000_000].numpy()
    t0 = time.time()
    carcs = og.random_access(sample, threads=threads)
    cpu_s = time.time() - t0
    alg = 8 * (k + 1) + 8 * arcs + 8 * k  # ids in, row pointers and successors out (the records read depend on the chains)
    del b_col, b_rp, d_ids
    return {"workload": f"successors(x) for {k} uniform random node ids (seed 3) on the headline graph, one batch", "queries": k, "arcs": int(arcs), "ms": ms,
            "nodes_per_s": k / (ms * 1e-3), "value": arcs / (ms * 1e-3), "unit": UNIT, "gpu_launches": int(launches), "verified": bool(ok),
            "verified_how": "every list equal to the row of the bulk-decoded CSR (itself checksum-verified)",
            "cpu_baseline": {"nodes_per_s": sample.size / cpu_s, "value": carcs / cpu_s, "unit": UNIT, "cores": threads, "kind": "port",
                             "sample": f"first {sample.size} of the same ids ({carcs} arcs) in {cpu_s:.2f} s, oracle successors(x) with recursive reference resolution (BVG:860-867, :1084) on {threads} threads"},
            "roofline": {"bound": "hbm", "achieved": alg / (ms * 1e-3) / 1e9, "unit": "GB/s", "frac": alg / (ms * 1e-3) / 1e9 / peaks()[0],
                         "algorithmic_bytes": alg, "note": "ids in + row pointers and successors out; bytes of the records read are not counted"}}


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return main_reference(args, rank, world)

    import torch
    import torch.distributed as dist

    import webgraph_big_b200 as wg

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the decode path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    threads = max(1, cpu_threads() // max(1, int(os.environ.get("LOCAL_WORLD_SIZE", world))))
    # one explicit stream for everything: the library launches on it and torch's events are recorded on it (the default
    # stream has handle 0, which the library takes as "create your own")
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)

    line = run_config(args, torch, dist, wg, rank, local_rank, world, dev, stream, threads, args.config)
    # BASELINE.json configs[4] under the same clock: at N=8 the 3e9-node graph fits the ranks' generation time budget
    if args.config == "c2" and world >= 8 and not args.no_configs:
        a5 = argparse.Namespace(**vars(args))
        a5.config = "c5"
        a5.steps, a5.warmup = max(3, min(args.steps, 5)), 3
        a5.no_e2e = True
        c5 = run_config(a5, torch, dist, wg, rank, local_rank, world, dev, stream, threads, "c5")
        if rank == 0:
            line["c5"] = c5
    if rank == 0:
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def run_config(args, torch, dist, wg, rank, local_rank, world, dev, stream, threads, config):
    if config == "c5":
        args.nodes = args.total_nodes // world
        args.mean_degree = 3.3
    sh = Shard(args, rank, world, threads)
    e, n, first = sh.e, sh.n, sh.first
    nodes_total = world * args.nodes
    t0 = time.time()
    # pinned copy of the compressed stream: the source of the per-step upload of the end-to-end measurement
    h_graph = torch.empty(sh.graph.size + 16, dtype=torch.uint8, pin_memory=True)
    h_graph[: sh.graph.size] = torch.from_numpy(sh.graph)
    g = wg.GpuBVGraph.load_slice(sh.properties_text(world, nodes_total), sh.slice_first, n + sh.halo, h_graph.numpy()[: sh.graph.size], sh.offsets_stream,
                                 halo_nodes=sh.halo, device=local_rank, cuda_stream=stream.cuda_stream)
    load_s = time.time() - t0

    d_rowptr = torch.empty(n + 1, dtype=torch.int64, device=dev)
    # arcs of the shard: the generator counts those of the whole part; the decode itself says how many the shard has
    arcs_cap = e.arcs
    d_col = torch.empty(max(arcs_cap, 1), dtype=torch.int64, device=dev)
    counts = torch.zeros(world, dtype=torch.int64, device=dev)
    mine = torch.zeros(1, dtype=torch.int64, device=dev)

    def step():
        g.decodeToCSRDevice(first, first + n, d_rowptr.data_ptr(), d_col.data_ptr(), arcs_cap, sync=False)
        if world > 1:
            # the only exchange of the path: per-shard arc counts -> global CSR offsets
            mine.copy_(d_rowptr[n:n + 1])
            dist.all_gather_into_tensor(counts, mine)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step()
    g.sync()
    stage_acc = np.zeros(5)
    launches0 = g.kernelLaunches()
    sampler = ClockSampler(local_rank)
    barrier()
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        step()
        stage_acc += np.array(g.lastStageMs())  # events already recorded by the library on this stream
    ev1.record()
    barrier()
    clocks = sampler.stop()
    g.sync()
    launches = g.kernelLaunches() - launches0
    fused_stats = g.fusedStats()
    ms = ev0.elapsed_time(ev1)
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = float(t.item()) / args.steps
    stage_ms = stage_acc / args.steps
    # per rank: device ms per step, halo rows decoded redundantly, nodes of the shard (rank 0 has no halo: the cost of the
    # halo path shows as the difference between its time and the others')
    mine_info = torch.tensor([ms / args.steps, float(sh.halo), float(n), float(stage_ms[4])], dtype=torch.float64, device=dev)
    info = torch.zeros(world * 4, dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_gather_into_tensor(info, mine_info)
    else:
        info.copy_(mine_info)
    info = info.cpu().reshape(world, 4).tolist()

    # parity at full size: the checksum of the decoded CSR must equal the one the generator took before encoding
    arcs = int(d_rowptr[n].item())
    chk = g.csrChecksumDevice(first, n, d_rowptr.data_ptr(), d_col.data_ptr())
    verified = chk == sh.checksum and (sh.lead > 0 or arcs == e.arcs)
    if not verified:
        raise SystemExit(f"rank {rank}: decoded CSR does not match the generator's checksum ({chk:#x} != {sh.checksum:#x}, {arcs} arcs)")
    global_base = int(counts[:rank].sum().item()) if world > 1 else 0

    full = config == "c2" and world == 1
    # the transpose of the CSR just decoded, device to device (what Transform.transposeOffline is for).
    # Checked at full size by the involution: the transpose of the transpose must hash to the generator's checksum.
    transpose = None
    if full and not args.no_transpose and n < (1 << 32):
        t_rp = torch.empty(n + 1, dtype=torch.int64, device=dev)
        t_col = torch.empty(max(arcs, 1), dtype=torch.int64, device=dev)
        g.transposeCSRDevice(n, d_rowptr.data_ptr(), d_col.data_ptr(), t_rp.data_ptr(), t_col.data_ptr())  # warm-up: sizes the scratch
        g.sync()
        reps = 3
        l0 = g.kernelLaunches()
        tv0, tv1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        tv0.record()
        for _ in range(reps):
            g.transposeCSRDevice(n, d_rowptr.data_ptr(), d_col.data_ptr(), t_rp.data_ptr(), t_col.data_ptr())
        tv1.record()
        g.sync()
        t_ms = tv0.elapsed_time(tv1) / reps
        t_launches = (g.kernelLaunches() - l0) // reps
        b_rp = torch.empty(n + 1, dtype=torch.int64, device=dev)
        g.transposeCSRDevice(n, t_rp.data_ptr(), t_col.data_ptr(), b_rp.data_ptr(), d_col.data_ptr())  # back into d_col
        g.sync()
        t_ok = g.csrChecksumDevice(first, n, b_rp.data_ptr(), d_col.data_ptr()) == e.checksum and bool(torch.equal(b_rp, d_rowptr))
        # algorithmic bytes: CSR in (8 B per arc + row pointers), transposed CSR out
        t_bytes = 2 * (8 * arcs + 8 * (n + 1))
        transpose = {"metric": "transposed arcs/s (CSR -> CSR, device resident)", "value": arcs / (t_ms * 1e-3), "unit": UNIT, "ms": t_ms,
                     "gpu_launches": int(t_launches), "verified_by_involution": bool(t_ok),
                     "roofline": {"bound": "hbm", "achieved": t_bytes / (t_ms * 1e-3) / 1e9, "unit": "GB/s",
                                  "frac": t_bytes / (t_ms * 1e-3) / 1e9 / peaks()[0], "algorithmic_bytes": t_bytes}}
        del t_rp, t_col, b_rp
        if not t_ok:
            raise SystemExit("transpose(transpose(CSR)) does not hash to the generator's checksum")

    # the device compressor on the CSR just decoded, with the generator's own compression splits: it must
    # write the very stream that was decoded, byte for byte (BVGraph.store -> .graph)
    encode = None
    if full and not args.no_configs:
        P = split_nodes(args.nodes)
        out = np.zeros(int(sh.graph.size) + 64, dtype=np.uint8)
        wg.encode_csr_device(min(n, 4 * P), d_rowptr.data_ptr(), d_col.data_ptr(), 0, max_ref_count=args.maxref, split_nodes=P, device=local_rank, out=out)  # warm-up
        t0 = time.perf_counter()
        enc, enc_bits, _ = wg.encode_csr_device(n, d_rowptr.data_ptr(), d_col.data_ptr(), arcs, max_ref_count=args.maxref, split_nodes=P, device=local_rank, out=out)
        enc_s = time.perf_counter() - t0
        e_ok = enc_bits == e.graph_bits and bool(np.array_equal(enc, e.graph[: enc.size]))
        encode = {"metric": "compressed arcs/s (device CSR -> .graph bytes in host memory)", "value": arcs / enc_s, "unit": UNIT, "ms": enc_s * 1e3,
                  "bits_per_arc": enc_bits / max(arcs, 1), "verified": bool(e_ok),
                  "verified_how": "the stream equals, byte for byte, the one the CPU restatement of the reference's compressor wrote (and that was decoded above)",
                  "host_encoder_s": sh.gen_s, "what": "wall time of wgbv_encode_csr: pricing of all (node, candidate) pairs, chain-length rule per split, emission, "
                                                      "copy of the stream to the host; scratch allocated inside the call"}
        del out
        if not e_ok:
            raise SystemExit("the device compressor did not reproduce the generator's stream")

    # configs[3] on the graph just decoded (the CSR in d_rowptr / d_col is the check)
    c4 = None
    if full and not args.no_configs:
        c4 = config_c4(torch, g, e, dev, d_rowptr, d_col, threads)

    # 10-bit labels on every arc of the graph just decoded (synthetic: arc j carries a hash of j), stored as a
    # BitStreamArcLabelledImmutableGraph stores them and decoded into an array parallel to the successors.
    labels = None
    if full and not args.no_labels:
        import bvgen
        width = 10
        t0 = time.time()
        h_rp = d_rowptr.cpu().numpy()
        lab, off = bvgen.write_labels_synthetic(width, h_rp)
        lab_gen_s = time.time() - t0
        L = wg.GpuArcLabels(wg.LABEL_FIXED_INT, width, lab, off, n, device=local_rank)
        d_lab = torch.empty(max(arcs, 1), dtype=torch.int64, device=dev)
        L.decodeDevice(g, first, first + n, d_rowptr.data_ptr(), d_lab.data_ptr(), arcs)
        g.sync()
        reps = 3
        lv0, lv1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        lv0.record()
        for _ in range(reps):
            L.decodeDevice(g, first, first + n, d_rowptr.data_ptr(), d_lab.data_ptr(), arcs)
        lv1.record()
        g.sync()
        l_ms = lv0.elapsed_time(lv1) / reps
        mult = bvgen.LABEL_HASH_MULTIPLIER - (1 << 64)  # the same multiplier as a signed 64-bit number
        l_ok = True
        for a in range(0, arcs, 1 << 27):
            j = torch.arange(a, min(arcs, a + (1 << 27)), dtype=torch.int64, device=dev)
            l_ok = l_ok and bool(torch.equal(((j * mult) >> 20) & ((1 << width) - 1), d_lab[a:a + j.numel()]))
            del j
        l_bytes = int(lab.size) + int(off.size) + 8 * arcs + 8 * (n + 1)
        labels = {"metric": "decoded arc labels/s (.labels -> int64 array parallel to the CSR)", "value": arcs / (l_ms * 1e-3), "unit": "labels/s",
                  "ms": l_ms, "label": f"FixedWidthIntLabel, {width} bits", "verified": bool(l_ok), "generation_s": lab_gen_s,
                  "roofline": {"bound": "hbm", "achieved": l_bytes / (l_ms * 1e-3) / 1e9, "unit": "GB/s",
                               "frac": l_bytes / (l_ms * 1e-3) / 1e9 / peaks()[0], "algorithmic_bytes": l_bytes}}
        L.close()
        del d_lab, lab, off, h_rp
        if not l_ok:
            raise SystemExit("decoded labels differ from the generator's")

    total_arcs = torch.tensor([arcs], dtype=torch.int64, device=dev)
    alg_t = torch.tensor([g.algorithmicBytes(first, first + n, arcs)], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(total_arcs)
        dist.all_reduce(alg_t)
    total_arcs = int(total_arcs.item())
    alg_bytes = int(alg_t.item())
    value = total_arcs / (ms_per_step * 1e-3)

    peak, peak_src = peaks()
    # the roofline is quoted for the whole step's device time (max over ranks) so that no stage is hidden: all ranks'
    # algorithmic bytes against world x the one-GPU peak
    achieved = alg_bytes / (ms_per_step * 1e-3) / 1e9
    # DRAM bytes actually moved per step, from the committed ncu capture of this very workload (None for other sizes)
    traffic = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "r02_traffic.json")))
        if config == "c2" and world == 1 and args.nodes == tr.get("nodes") and args.maxref == 3 and args.max_degree == 10000:
            traffic = tr["dram_bytes_per_step"]
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak * world, "unit": "GB/s", "frac": achieved / (peak * world), "traffic": traffic,
                "traffic_source": "profiles/r02_traffic.json (ncu capture of this workload on the committed build)" if traffic else None,
                "peak_source": peak_src + (f" x {world} GPUs" if world > 1 else ""), "frac_of_nominal_8TBs": achieved / (8000.0 * world),
                "kernel": "whole decode step (all launches)", "algorithmic_bytes_per_step": alg_bytes,
                "stage_ms_rank0": {"header": float(stage_ms[0]), "scan": float(stage_ms[1]), "depth": float(stage_ms[2]), "successors": float(stage_ms[3]),
                                   "total": float(stage_ms[4])}}

    # end to end through the C ABI with HOST buffers: upload of the compressed stream + decode + read-back of the
    # CSR, chunk by chunk into pinned buffers that are reused (what a host-side NodeIterator consumes)
    def run_e2e(narrow):
        chunk = 1 << 22
        d2h = 0
        esz = 4 if narrow else 8
        consumed = np.zeros(1, dtype=np.int64)
        for a, b, rp, col in g.scan(first, first + min(n, 3 * chunk + 1), chunk, narrow=narrow):  # untimed warm-up: the library sizes its chunk buffers
            pass
        barrier()
        t0 = time.perf_counter()
        for _ in range(max(1, args.e2e_steps)):
            got = 0
            stamps = [time.perf_counter()]
            # the streaming form of nodeIterator(): the compressed stream goes up piece by piece from pinned host memory,
            # chunks come back in pinned host buffers owned by the library while the following chunks are decoded
            for a, b, rp, col in g.scan(first, first + n, chunk, upload=(h_graph.data_ptr(), sh.graph.size), narrow=narrow):
                got += col.size
                d2h += (b - a + 1) * 8 + col.size * esz
                consumed[0] += int(rp[-1]) + (int(col[-1]) if col.size else 0)  # touch the result on the host
                stamps.append(time.perf_counter())
            assert got == arcs
        if args.e2e_trace and rank == 0:
            print("e2e chunk arrivals (ms since the previous one):", " ".join(f"{(y - x) * 1e3:.1f}" for x, y in zip(stamps, stamps[1:])),
                  file=sys.stderr)
        barrier()
        dt = (time.perf_counter() - t0) / max(1, args.e2e_steps)
        tt = torch.tensor([dt], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return {"value": total_arcs / float(tt.item()), "unit": UNIT, "h2d_bytes_per_step": int(sh.graph.size) * world,
                "d2h_bytes_per_step": int(d2h // max(1, args.e2e_steps)) * world, "ms_per_step": float(tt.item()) * 1e3,
                "successor_type": "uint32" if narrow else "int64",
                "how": f"wgbv_scan_begin_upload / wgbv_scan_next over {chunk}-node chunks: upload of the compressed stream from pinned "
                       "host memory, decode on the device and copy-back to pinned host buffers, the three overlapped chunk by chunk"}

    # e2e: successors delivered as int64, the reference's own type (long).  e2e_narrow: the same scan with
    # wgbv_scan_set_narrow (uint32 successors, same values, half the PCIe bytes), reported beside it, never instead of it.
    e2e = e2e_narrow = None
    if not args.no_e2e:
        e2e = run_e2e(False)
        if nodes_total <= 1 << 32:
            e2e_narrow = run_e2e(True)

    pcie = None
    if not args.no_e2e and config == "c2":
        pcie = pcie_probe(torch, dist, dev, world)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        th = cpu_threads()
        sample = args.cpu_sample_nodes or min(n, 1_000_000 * th)
        rate, carcs, cdt = run_cpu(e, sample, th)
        cpu = {"value": rate, "unit": UNIT, "cores": th, "kind": "port",
               "sample": f"first {sample} of {n} nodes ({carcs} arcs) in {cdt:.2f} s, oracle port of nodeIterator split over {th} threads"}

    c1 = c3 = None
    if full and not args.no_configs:
        c1 = config_c1(torch, wg, dev, local_rank, stream)
        g.close()
        g = None
        c3 = config_c3(torch, wg, args, dev, local_rank, stream, d_rowptr, d_col, n, arcs, e.checksum, threads)

    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64",
            "data": "synthetic",
            "config": {"workload": workload_name(args, world), "nodes_per_gpu": args.nodes, "nodes_rank0": n, "arcs_rank0": arcs, "total_arcs": total_arcs,
                       "compressed_bytes_rank0": int(sh.graph.size), "bits_per_arc": e.graph_bits / max(e.arcs, 1),
                       "encoder_statistics_rank0": stats_of(e),
                       "l2": "inputs and outputs exceed L2 (compressed stream and 8 B/arc output are GBs per step)",
                       "parallelism": (f"{world} contiguous shards of one graph, boundaries inside compression splits; every rank but the first decodes the "
                                       f"{int(max(r[1] for r in info))} rows before its shard as halo (no data-path collective); all-gather of arc counts") if world > 1
                       else "1 shard; no collective",
                       "halo_nodes": int(max(r[1] for r in info)), "halo_nodes_by_rank": [int(r[1]) for r in info],
                       "ms_per_step_by_rank": [r[0] for r in info], "decode_ms_by_rank": [r[3] for r in info],
                       "decode_ms_by_rank_what": "the library's own events around the decode of the shard, without the all-gather: rank 0 has no halo (one-launch front end), the others decode 21 halo rows (separate header and scan launches)",
                       "nodes_by_rank": [int(r[2]) for r in info]},
            "roofline": roofline, "e2e": e2e, "e2e_narrow": e2e_narrow, "pcie_probe": pcie, "encode": encode, "transpose": transpose, "labels": labels, "c1": c1, "c3": c3, "c4": c4,
            "cpu_baseline": cpu, "gpu_launches": int(launches), "clocks": clocks,
            "verified": bool(verified), "fused": fused_stats, "global_row_base_rank0": global_base, "generation_s": sh.gen_s, "load_s": load_s,
        }
    if g is not None:
        g.close()
    del d_rowptr, d_col, h_graph
    return line


if __name__ == "__main__":
    sys.exit(main())
