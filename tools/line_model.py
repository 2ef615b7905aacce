#!/usr/bin/env python
"""L1 line model for the lj/cut gather (no GPU needed).  ubench2 (profiles/r02_ubench.md) shows the L1 data pipe spends one
wavefront per distinct 128-byte LINE of a warp-wide LDG.256 gather (4 position records per line), not one per sector.
This script counts distinct lines per warp step of the per-atom kernel (lane = atom, step k = k-th list entry, lists in
(dz,dy)-row order then ascending index, as k_build_bins emits them) on a thermalised melt, for several within-bin sort
orders.  Usage: python tools/line_model.py [cells]"""
import sys
from pathlib import Path

import numpy as np
from scipy.spatial import cKDTree

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import oracle  # noqa: E402  (development tool, not product)

cells = int(sys.argv[1]) if len(sys.argv) > 1 else 20
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
nb = int(L / 2.8)
bs = L / nb
b = np.minimum((x / bs).astype(int), nb - 1)
binkey = (b[:, 2] * nb + b[:, 1]) * nb + b[:, 0]
rng = np.random.default_rng(1)


def fine_key(sub, kind):
    f = np.minimum(((x - b * bs) / (bs / sub)).astype(int), sub - 1)
    if kind == "morton":
        k = np.zeros(n, dtype=np.int64)
        for bit in range(3):
            for d in range(3):
                k |= ((f[:, d] >> bit) & 1).astype(np.int64) << (3 * bit + d)
        return k
    if kind == "x":
        return ((x[:, 0] - b[:, 0] * bs) / bs * 4096).astype(np.int64)
    if kind == "zyx":
        return (f[:, 2] * sub + f[:, 1]) * sub + f[:, 0]
    if kind == "snake":  # boustrophedon through the sub-cells
        fy = np.where(f[:, 2] % 2 == 0, f[:, 1], sub - 1 - f[:, 1])
        fx = np.where((f[:, 2] * sub + fy) % 2 == 0, f[:, 0], sub - 1 - f[:, 0])
        return (f[:, 2] * sub + fy) * sub + fx
    return rng.integers(0, 1 << 30, n)


def evaluate(name, within):
    perm = np.lexsort((within, binkey))
    rank = np.empty(n, dtype=np.int64)
    rank[perm] = np.arange(n)
    ri, rj = rank[I], rank[J]
    # row of j relative to i's bin (periodic), rows ordered (dz, dy); inside a row ascending index
    dy = (b[J, 1] - b[I, 1] + nb + 1) % nb
    dz = (b[J, 2] - b[I, 2] + nb + 1) % nb
    # periodic wrap in x changes index order inside a row; use unwrapped bin x for the order
    dxb = (b[J, 0] - b[I, 0] + nb + 1) % nb
    order = np.lexsort((rj, dxb, dy, dz, ri))
    ri, rj = ri[order], rj[order]
    start = np.searchsorted(ri, np.arange(n))
    cnt = np.diff(np.append(start, len(ri)))
    k = np.arange(len(ri)) - start[ri]
    warp = ri // 32
    key = warp * 256 + k
    # distinct lines and sectors per (warp, step)
    line = rj // 4
    u = np.unique(np.stack([key, line], 1), axis=0)
    nl = np.bincount(u[:, 0])
    nl = nl[nl > 0]
    u2 = np.unique(np.stack([key, rj], 1), axis=0)
    ns = np.bincount(u2[:, 0])
    ns = ns[ns > 0]
    lanes = np.bincount(key)
    lanes = lanes[lanes > 0]
    nw = (n + 31) // 32
    steps = np.maximum.reduceat(cnt, np.arange(0, n, 32))
    print(f"{name:28s}: steps/warp {steps.mean():6.1f}, active lanes/step {lanes.mean():5.1f}, distinct sectors/step {ns.mean():5.1f}, "
          f"distinct 128B lines/step {nl.mean():5.1f}  -> lines per warp {nl.sum() / nw:7.0f}")


evaluate("within bin: random", fine_key(1, "rand"))
evaluate("within bin: x", fine_key(1, "x"))
for sub in (2, 3, 4):
    evaluate(f"within bin: zyx cells /{sub}", fine_key(sub, "zyx"))
    evaluate(f"within bin: snake /{sub}", fine_key(sub, "snake"))
evaluate("within bin: morton /4", fine_key(4, "morton"))
evaluate("within bin: morton /8", fine_key(8, "morton"))


def evaluate_rows(name, within, nrows_merge=1):
    """row-aligned lists: every lane of a warp walks the same (dz,dy) row at the same time; a row takes as many steps
    as its longest sub-list in the warp (shorter ones are padded)"""
    perm = np.lexsort((within, binkey))
    rank = np.empty(n, dtype=np.int64)
    rank[perm] = np.arange(n)
    ri, rj = rank[I], rank[J]
    dy = (b[J, 1] - b[I, 1] + nb + 1) % nb
    dz = (b[J, 2] - b[I, 2] + nb + 1) % nb
    dxb = (b[J, 0] - b[I, 0] + nb + 1) % nb
    row = (dz * 3 + dy) // nrows_merge
    order = np.lexsort((rj, dxb, row, ri))
    ri, rj, row = ri[order], rj[order], row[order]
    # position inside (atom, row)
    ar = ri * 16 + row
    first = np.r_[True, ar[1:] != ar[:-1]]
    startpos = np.maximum.accumulate(np.where(first, np.arange(len(ar)), 0))
    k = np.arange(len(ar)) - startpos
    warp = ri // 32
    key = (warp * 16 + row) * 64 + k
    line = rj // 4
    u = np.unique(np.stack([key, line], 1), axis=0)
    nl = np.bincount(u[:, 0])
    nl = nl[nl > 0]
    nsteps = len(nl)
    nw = (n + 31) // 32
    print(f"{name:28s}: row-aligned (merge {nrows_merge}): steps/warp {nsteps / nw:6.1f}, lanes/step {len(ar) / nsteps:5.1f}, lines/step {nl.mean():5.1f} "
          f"-> lines per warp {nl.sum() / nw:7.0f}, DP cost {nsteps / nw * 8.5:6.0f} cyc")


for m in (1, 3):
    evaluate_rows("within bin: random", fine_key(1, "rand"), m)
    evaluate_rows("within bin: x", fine_key(1, "x"), m)
    evaluate_rows("within bin: morton /4", fine_key(4, "morton"), m)
    evaluate_rows("within bin: zyx /2", fine_key(2, "zyx"), m)
