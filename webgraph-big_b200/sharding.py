"""Multi-GPU plumbing of the decode path: one process per GPU, contiguous bit-balanced shards, and the only exchange
the path has -- an all-gather of per-shard arc counts that turns local CSR row pointers into global ones.

torch.distributed is used for the collective (NCCL on GPUs, gloo in the CPU tests); nothing here touches the decode
kernels.  The shard plan is the library's own (wgbv_plan_shards, host only).
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import Props


def plan_shards(props, offsets_bytes, shard_count):
    """(bounds, halo): bounds[k] = first node of shard k, bounds[shard_count] = nodes; halo = window * maxref."""
    if isinstance(props, str):
        p = Props()
        _lib.check(_lib.lib().wgbv_parse_properties(props.encode("utf-8"), C.byref(p)))
        props = p
    o = np.ascontiguousarray(np.frombuffer(offsets_bytes, dtype=np.uint8) if not isinstance(offsets_bytes, np.ndarray) else offsets_bytes)
    bounds = np.zeros(shard_count + 1, dtype=np.int64)
    halo = C.c_int64()
    _lib.check(_lib.lib().wgbv_plan_shards(C.byref(props), o.ctypes.data, o.size, shard_count, bounds.ctypes.data, C.byref(halo)))
    return bounds, halo.value


def global_row_base(local_arcs, group=None, device=None):
    """All-gathers the arc counts of the shards; returns (base of this rank's rows in the global CSR, total arcs,
    all counts).  With no process group initialised it is the single-shard identity."""
    import torch
    import torch.distributed as dist

    if not dist.is_available() or not dist.is_initialized():
        return 0, int(local_arcs), [int(local_arcs)]
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    mine = torch.tensor([int(local_arcs)], dtype=torch.int64, device=device)
    counts = torch.zeros(world, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(counts, mine, group=group)
    counts = [int(c) for c in counts.tolist()]
    return sum(counts[:rank]), sum(counts), counts
