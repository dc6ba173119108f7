"""The shards bench.py hands to the ranks of a multi-GPU run are pieces of ONE graph: every rank generates the
compression splits that cover its shard, and the rows it decodes -- and their checksum -- are those of the graph
generated in one piece.  CPU only (generator + oracle)."""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402
import bvgen  # noqa: E402
import bvoracle  # noqa: E402


def _args(nodes):
    return argparse.Namespace(nodes=nodes, seed=2, mean_degree=12.0, maxref=3, max_degree=10000, config="c2", giant_every=0, giant_degree=0)


def _decode(e):
    p = bvoracle.params_from_properties(bvoracle.parse_properties(e.properties_text()))
    p.nodes = e.n
    deg, col = bvoracle.OracleGraph(p, e.graph, offsets=e.offsets).decode_range()
    rp = np.zeros(deg.size + 1, dtype=np.int64)
    np.cumsum(deg, out=rp[1:])
    return rp, col


def test_shards_are_pieces_of_one_graph():
    n, world = 30000, 3
    args = _args(n)
    P = bench.split_nodes(n)
    whole = bvgen.copy_model(world * n, seed=args.seed, mean_degree=args.mean_degree, nsplits=-P, global_n=world * n)
    wrp, wcol = _decode(whole)
    bounds = bench.shard_bounds(args, world)
    assert bounds[0] == 0 and bounds[-1] == world * n and all(b % P == P // 2 for b in bounds[1:-1])
    used_halo = 0
    for rank in range(world):
        sh = bench.Shard(args, rank, world, threads=2)
        assert (sh.first, sh.last) == (bounds[rank], bounds[rank + 1])
        rp, col = _decode(sh.e)
        col = col + sh.part_first  # (the oracle numbered the part's rows from 0; successors are stored relative to the row)
        lead = sh.lead
        assert np.array_equal(col[rp[lead]:], wcol[wrp[sh.first]:wrp[sh.last]])
        assert np.array_equal(rp[lead:] - rp[lead], wrp[sh.first:sh.last + 1] - wrp[sh.first])
        assert sh.checksum == bvgen.csr_checksum(sh.first, wrp[sh.first:sh.last + 1] - wrp[sh.first], wcol[wrp[sh.first]:wrp[sh.last]])
        if rank:
            assert sh.halo == 21 and sh.slice_first == sh.first - 21
            # the slice handed to the library: its offsets stream gives the records' positions inside the cut bit stream
            p = bvoracle.params_from_properties(bvoracle.parse_properties(sh.e.properties_text()))
            p.nodes = sh.n + sh.halo
            og = bvoracle.OracleGraph(p, sh.graph, sh.offsets_stream.tobytes())
            deg, c2 = og.decode_range(sh.halo, sh.n + sh.halo)
            # (ids are global in the lists; the oracle adds gaps to local row numbers, so compare differences to the row id)
            want = wcol[wrp[sh.first]:wrp[sh.last]] - np.repeat(np.arange(sh.first, sh.last), np.diff(wrp[sh.first:sh.last + 1]))
            got = c2 - np.repeat(np.arange(sh.halo, sh.n + sh.halo), deg)
            assert np.array_equal(got, want)
            used_halo += 1
    assert used_halo == world - 1
