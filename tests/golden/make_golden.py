"""Regenerates the committed golden fixtures from the reference checkout (run in the build container only).

Inputs (read-only, reference repo): slow/it/unimi/dsi/big/webgraph/cnr-2000.{graph,offsets,properties,graph-txt.gz}
-- the known-answer fixture behind the reference's BVGraphTest.testLarge (test/.../BVGraphTest.java:105-123).

Outputs (committed under tests/golden/):
  cnr-2000.graph / .offsets / .properties   the compressed triple, byte-for-byte (it is input DATA, not code)
  cnr-2000.golden.npz                       the ASCII golden (ASCIIGraph format, ASCIIGraph.java:57-61: first line n,
                                            then one line of increasing successors per node) parsed into
                                            deg[int32 n] and delta-coded successors col_delta[int32 m]
                                            (first successor of each row stored as is, then gaps)
Nothing at test time reads /root/reference: the GPU box does not have it.
"""
import gzip
import os
import shutil
import sys

import numpy as np

REF = "/root/reference/slow/it/unimi/dsi/big/webgraph"
HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    for ext in ("graph", "offsets", "properties"):
        shutil.copyfile(f"{REF}/cnr-2000.{ext}", f"{HERE}/cnr-2000.{ext}")
        os.chmod(f"{HERE}/cnr-2000.{ext}", 0o644)
    with gzip.open(f"{REF}/cnr-2000.graph-txt.gz", "rt") as f:
        n = int(f.readline())
        deg = np.zeros(n, dtype=np.int32)
        chunks = []
        for i in range(n):
            row = np.array(f.readline().split(), dtype=np.int64)
            deg[i] = row.size
            if row.size:
                assert (np.diff(row) > 0).all(), i
                d = row.copy()
                d[1:] = np.diff(row)
                chunks.append(d)
        assert f.readline() == ""
    col_delta = np.concatenate(chunks)
    assert col_delta.max() < 2**31
    np.savez_compressed(f"{HERE}/cnr-2000.golden.npz", deg=deg, col_delta=col_delta.astype(np.int32))
    print(n, int(deg.sum()), os.path.getsize(f"{HERE}/cnr-2000.golden.npz"))


if __name__ == "__main__":
    sys.exit(main())
