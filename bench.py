#!/usr/bin/env python
"""bench.py -- decoded arcs/s of the BVGraph -> CSR hot path (BASELINE.json metric).

A "step" is one complete bulk decode of the rank's graph: compressed bit stream + Elias-Fano offsets resident in
HBM -> int64 CSR row pointers and successors resident in HBM.  The workload at N=1 is BASELINE.json configs[1]:
a synthetic copy-model web-like graph, 100M nodes / ~2B arcs, default BVGraph parameters (w=7, maxRef=3,
minInt=4, zeta_3), produced by this repository's generator + encoder (datagen/).  At N>1 every rank holds one
such part of a single N x 100M-node graph (global node ids; parts are cut at compression-split boundaries, so no
halo is needed) and the ranks all-gather their arc counts over NCCL to form global CSR offsets: weak scaling.

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
    return (f"copy-model web-like graph, {args.nodes // 1_000_000}M nodes/GPU x {world} GPU(s), mean degree {args.mean_degree:g}, "
            f"BVGraph w=7 maxRef={args.maxref} minInt=4 zeta3, bulk decode to int64 CSR")


def generate(args, rank, world, threads):
    import bvgen
    t0 = time.time()
    e = bvgen.copy_model(args.nodes, seed=args.seed, mean_degree=args.mean_degree, params=bvgen.make_params(maxref=args.maxref), max_degree=args.max_degree,
                         nsplits=256, threads=threads, first_node=rank * args.nodes, global_n=world * args.nodes)
    return e, time.time() - t0


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


def run_cpu(e, sample_nodes, threads, repeats=1):
    """The reference's CPU algorithm (oracle port of BVGraph.nodeIterator / splitNodeIterators) over the first
    `sample_nodes` nodes, all host threads; returns (arcs/s, arcs, seconds)."""
    import bvoracle
    p = bvoracle.params_from_properties(bvoracle.parse_properties(e.properties_text()))
    # parts of a larger graph are decoded with local ids: the oracle only needs consistent offsets
    p.nodes = e.n
    og = bvoracle.OracleGraph(p, e.graph, offsets=e.offsets)
    best = None
    arcs = 0
    for _ in range(repeats):
        t0 = time.time()
        deg, _ = og.decode_range(0, sample_nodes, threads=threads, want_col=False)
        dt = time.time() - t0
        arcs = int(deg.sum())
        best = dt if best is None else min(best, dt)
    return arcs / best, arcs, best


def main_reference(args, rank, world):
    """--impl reference: the reference's own algorithm on the host cores (no JVM on this image: the C restatement
    in oracle/, pinned on the reference's cnr-2000 fixture, run splitNodeIterators-style on all cores)."""
    if rank != 0:
        return 0
    threads = cpu_threads()
    e, gen_s = generate(args, 0, 1, threads)
    sample = args.cpu_sample_nodes or min(e.n, 1_000_000 * threads)
    for _ in range(max(args.warmup, 0)):
        run_cpu(e, min(sample, 200_000 * threads), threads)
    times, arcs = [], 0
    for _ in range(max(args.steps, 1)):
        rate, arcs, dt = run_cpu(e, sample, threads)
        times.append(dt)
    dt = float(np.mean(times))
    value = arcs / dt
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "config": {"workload": workload_name(args, 1), "sample": f"first {sample} nodes ({arcs} arcs) per step"},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"first {sample} of {e.n} nodes ({arcs} arcs), splitNodeIterators-style over {threads} threads, degrees + successor lists decoded"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "generation_s": gen_s,
    }
    print(json.dumps(line))
    return 0


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
    e, gen_s = generate(args, rank, world, threads)
    n, arcs = e.n, e.arcs
    first = rank * n

    props_text = e.properties_text().replace(f"nodes={n}", f"nodes={world * n}")
    t0 = time.time()
    # one explicit stream for everything: the library launches on it and torch's events are recorded on it (the default
    # stream has handle 0, which the library takes as "create your own")
    stream = torch.cuda.Stream(device=dev)
    torch.cuda.set_stream(stream)
    # pinned copy of the compressed stream: the source of the per-step upload of the end-to-end measurement
    h_graph = torch.empty(e.graph.size + 16, dtype=torch.uint8, pin_memory=True)
    h_graph[: e.graph.size] = torch.from_numpy(e.graph)
    g = wg.GpuBVGraph.load_slice(props_text, first, n, h_graph.numpy()[: e.graph.size], e.offsets_stream, device=local_rank,
                                 cuda_stream=stream.cuda_stream)
    load_s = time.time() - t0

    d_rowptr = torch.empty(n + 1, dtype=torch.int64, device=dev)
    d_col = torch.empty(arcs, dtype=torch.int64, device=dev)
    counts = torch.zeros(world, dtype=torch.int64, device=dev)
    mine = torch.zeros(1, dtype=torch.int64, device=dev)

    def step():
        g.decodeToCSRDevice(first, first + n, d_rowptr.data_ptr(), d_col.data_ptr(), arcs, sync=False)
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

    # parity at full size: the checksum of the decoded CSR must equal the one the generator took before encoding
    chk = g.csrChecksumDevice(first, n, d_rowptr.data_ptr(), d_col.data_ptr())
    verified = (chk == e.checksum) and int(d_rowptr[n].item()) == arcs
    if not verified:
        raise SystemExit(f"rank {rank}: decoded CSR does not match the generator's checksum ({chk:#x} != {e.checksum:#x})")
    global_base = int(counts[:rank].sum().item()) if world > 1 else 0

    # the transpose of the CSR just decoded, device to device (what Transform.transposeOffline is for).
    # Checked at full size by the involution: the transpose of the transpose must hash to the generator's checksum.
    transpose = None
    if world == 1 and not args.no_transpose and n < (1 << 32):
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

    # 10-bit labels on every arc of the graph just decoded (synthetic: arc j carries a hash of j), stored as a
    # BitStreamArcLabelledImmutableGraph stores them and decoded into an array parallel to the successors.
    labels = None
    if world == 1 and not args.no_labels:
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
    if world > 1:
        dist.all_reduce(total_arcs)
    total_arcs = int(total_arcs.item())
    value = total_arcs / (ms_per_step * 1e-3)

    alg_bytes = g.algorithmicBytes(first, first + n, arcs)
    peak, peak_src = peaks()
    # the dominant kernel group is the successor decode (stage 3); the roofline is quoted for the whole step's
    # device time so that no stage is hidden
    dom_ms = float(stage_ms[4])
    achieved = alg_bytes / (dom_ms * 1e-3) / 1e9
    # DRAM bytes actually moved per step, from the committed ncu capture of this very workload (None for other sizes)
    traffic = None
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json")))
        if args.nodes == 100_000_000 and args.maxref == 3 and args.max_degree == 10000:
            traffic = tr["dram_bytes_per_step"]
    except Exception:
        pass
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                "peak_source": peak_src, "kernel": "whole decode (header + scan + depth + successor decode)",
                "algorithmic_bytes_per_step": alg_bytes, "stage_ms": {"header": float(stage_ms[0]), "scan": float(stage_ms[1]),
                                                                     "depth": float(stage_ms[2]), "successors": float(stage_ms[3]),
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
            for a, b, rp, col in g.scan(first, first + n, chunk, upload=(h_graph.data_ptr(), e.graph.size), narrow=narrow):
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
        return {"value": total_arcs / float(tt.item()), "unit": UNIT, "h2d_bytes_per_step": int(e.graph.size) * world,
                "d2h_bytes_per_step": int(d2h // max(1, args.e2e_steps)) * world, "ms_per_step": float(tt.item()) * 1e3,
                "successor_type": "uint32" if narrow else "int64",
                "how": f"wgbv_scan_begin_upload / wgbv_scan_next over {chunk}-node chunks: upload of the compressed stream from pinned "
                       "host memory, decode on the device and copy-back to pinned host buffers, the three overlapped chunk by chunk"}

    # e2e: successors delivered as int64, the reference's own type (long).  e2e_narrow: the same scan with
    # wgbv_scan_set_narrow (uint32 successors, same values, half the PCIe bytes), reported beside it, never instead of it.
    e2e = e2e_narrow = None
    if not args.no_e2e:
        e2e = run_e2e(False)
        if n * world <= 1 << 32:
            e2e_narrow = run_e2e(True)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        th = cpu_threads()
        sample = args.cpu_sample_nodes or min(n, 1_000_000 * th)
        rate, carcs, cdt = run_cpu(e, sample, th)
        cpu = {"value": rate, "unit": UNIT, "cores": th, "kind": "port",
               "sample": f"first {sample} of {n} nodes ({carcs} arcs) in {cdt:.2f} s, oracle port of nodeIterator split over {th} threads"}

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int64",
            "data": "synthetic",
            "config": {"workload": workload_name(args, world), "nodes_per_gpu": n, "arcs_per_gpu_rank0": arcs, "total_arcs": total_arcs,
                       "compressed_bytes_rank0": int(e.graph.size), "bits_per_arc": e.graph_bits / max(arcs, 1),
                       "l2": "inputs and outputs exceed L2 (compressed stream and 8 B/arc output are GBs per step)",
                       "parallelism": f"{world} contiguous shard(s), no data-path collective; all-gather of arc counts"},
            "roofline": roofline, "e2e": e2e, "e2e_narrow": e2e_narrow, "transpose": transpose, "labels": labels, "cpu_baseline": cpu, "gpu_launches": int(launches), "clocks": clocks,
            "verified": bool(verified), "fused": fused_stats, "global_row_base_rank0": global_base, "generation_s": gen_s, "load_s": load_s,
        }
        print(json.dumps(line))
    g.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
