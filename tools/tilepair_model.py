#!/usr/bin/env python
"""Register-tile model for the lj/cut kernel (no GPU needed).  A warp owns a tile of 32 consecutive atoms (lane = atom i)
and walks the j-tiles its atoms have neighbors in; the 32 atoms of a j-tile sit one per lane in registers and a lane
fetches its s-th neighbor of that tile with an indexed shuffle.  Steps per (i-tile, j-tile) = the largest number of
neighbors any lane has in that j-tile, so the FP64 lane efficiency is pairs / (32 * steps).  The script measures that
efficiency on a thermalised LJ melt for several sort orders.
Usage: python tools/tilepair_model.py [cells] [sortcell] [tile]"""
import sys
from pathlib import Path

import numpy as np
from scipy.spatial import cKDTree

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import oracle  # noqa: E402  (development tool, not product)

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 20
sortcell = float(sys.argv[2]) if len(sys.argv) > 2 else 2.8
TILE = int(sys.argv[3]) if len(sys.argv) > 3 else 32
s = oracle.make_lj_system(cells)
s.setup()
s.run(100)
x = s.atoms()["x"].copy()
L = cells * (4 / 0.8442) ** (1 / 3)
x %= L
n = len(x)
tree = cKDTree(x, boxsize=L)
pr = tree.query_pairs(2.8, output_type="ndarray")
I = np.concatenate([pr[:, 0], pr[:, 1]])
J = np.concatenate([pr[:, 1], pr[:, 0]])
print(f"{n} atoms, {len(I) / n:.1f} neighbors per atom (full)")


def order(kind, c):
    nb = int(L / c)
    b = np.minimum((x / (L / nb)).astype(int), nb - 1)
    if kind == "row":  # x fastest
        key = (b[:, 2] * nb + b[:, 1]) * nb + b[:, 0]
    elif kind == "morton":
        key = np.zeros(n, dtype=np.int64)
        for bit in range(8):
            for d in range(3):
                key |= ((b[:, d] >> bit) & 1).astype(np.int64) << (3 * bit + d)
    elif kind == "block":  # 2x2x2 blocks of cells, then x fastest inside
        B = b // 2
        nB = (nb + 1) // 2
        key = (((B[:, 2] * nB + B[:, 1]) * nB + B[:, 0]) * 8 + (b[:, 2] % 2) * 4 + (b[:, 1] % 2) * 2 + (b[:, 0] % 2))
    return np.argsort(key, kind="stable")


for kind, c in (("row", 2.8), ("row", sortcell), ("morton", sortcell), ("block", sortcell), ("morton", sortcell / 2)):
    perm = order(kind, c)
    rank = np.empty(n, dtype=np.int64)
    rank[perm] = np.arange(n)
    ti, tj = rank[I] // TILE, rank[J] // TILE
    ntile = (n + TILE - 1) // TILE
    # neighbors of atom i inside j-tile tj
    key = rank[I] * ntile + tj
    u, cnt = np.unique(key, return_counts=True)
    ui, utj = u // ntile, u % ntile
    tkey = (ui // TILE) * ntile + utj
    o = np.argsort(tkey, kind="stable")
    tk, start = np.unique(tkey[o], return_index=True)
    mx = np.maximum.reduceat(cnt[o], start)
    steps = mx.sum()
    npairs_tiles = len(tk)
    eff = len(I) / (TILE * steps)
    # half: only j-tiles >= i-tile (Newton-3 across tiles would need f_j reductions; here for reference)
    print(f"sort {kind:6s} cell {c:4.2f}: tile pairs per i-tile {npairs_tiles / ntile:6.1f}, steps per i-tile {steps / ntile:7.1f} "
          f"(ideal {len(I) / n:5.1f}), lane efficiency {eff:5.3f}, mean max-count {mx.mean():.2f}")
