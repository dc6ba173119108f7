"""The N > 1 path on CPU: world_size-2 and -3 gloo process groups run the host-side logic of the sharded decode --
the library's bit-balanced shard plan, per-shard decoding with the window*maxref halo (done here by the CPU oracle,
since there is no GPU), and the all-gather of arc counts that forms the global CSR offsets."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, basename, q):
    for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.join(ROOT, "datagen"), os.path.join(ROOT, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist

    import bvoracle as bo
    import webgraph_big_b200  # noqa: F401
    from webgraph_big_b200 import sharding

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        og = bo.OracleGraph.load(basename)
        props = open(basename + ".properties").read()
        bounds, halo = sharding.plan_shards(props, open(basename + ".offsets", "rb").read(), world)
        lo, hi = int(bounds[rank]), int(bounds[rank + 1])
        deg, col = og.decode_range(lo, hi)  # nodeIterator(lo): pre-fills the window from the halo
        base, total, counts = sharding.global_row_base(col.size)
        bits = int(og.offsets[hi] - og.offsets[lo])
        q.put((rank, lo, hi, halo, base, total, counts, bits, deg, col))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_decode_plumbing_over_gloo(world, cnr_basename, cnr_golden):
    import torch.multiprocessing as mp

    deg, rowptr, col = cnr_golden
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000 + world
    procs = [ctx.Process(target=_worker, args=(r, world, port, cnr_basename, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = sorted((q.get(timeout=180) for _ in range(world)), key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    nxt = 0
    for rank, lo, hi, halo, base, total, counts, bits, d, c in got:
        assert lo == nxt and halo == 21
        nxt = hi
        assert base == rowptr[lo] and total == col.size and counts[rank] == c.size
        assert np.array_equal(d, deg[lo:hi]) and np.array_equal(c, col[rowptr[lo]:rowptr[hi]])
        # global row pointers of the shard = local ones + the gathered base
        local_rp = np.zeros(d.size + 1, dtype=np.int64)
        np.cumsum(d, out=local_rp[1:])
        assert np.array_equal(local_rp + base, rowptr[lo:hi + 1])
    assert nxt == deg.size
    # the plan balances compressed bits, not node counts
    bits = [t[7] for t in got]
    assert max(bits) - min(bits) < 0.02 * sum(bits)


def test_single_process_identity():
    from webgraph_big_b200 import sharding
    assert sharding.global_row_base(123) == (0, 123, [123])
