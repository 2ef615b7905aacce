#!/usr/bin/env python
"""Shared-memory wavefront model for the tiled lj/cut kernel (no GPU needed): reconstructs, for a thermalised LJ melt,
the staging layout of interior bin segments and the slot lists of their atoms as k_build_bins / k_tile_convert produce
them (rows in (dz, dy) order, slots ascending), then counts per warp step
  * distinct 32-byte sectors   -- the L1 gather of the default kernel (ncu measured 26 wavefronts per warp step),
  * LDS wavefronts for 32-byte records (one LDS.128 in 4 quarter-warp phases + one LDS.64 in 2 half-warp phases),
  * LDS wavefronts for the x[] y[] z[] layout (three LDS.64, 2 half-warp phases each),
a wavefront per phase being the largest number of DISTINCT addresses that fall into one bank (same address = broadcast).
Usage: python tools/tiled_bank_model.py [cells] [G]      (positions come from an emulated 60-step run)"""
import ctypes as C
import importlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path[:0] = [str(ROOT), str(ROOT / "tests"), str(ROOT / "tests" / "cudaemu")]


def melt(cells):
    import build as emu_build
    pkg = importlib.import_module("lammps-fugaku-threadpool_b200")
    L = C.CDLL(str(emu_build.build()))
    for name, (res, args) in pkg.lib.PROTOTYPES.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    pkg.lib._lib, pkg.lib.load = L, (lambda rebuild=False: L)
    inp = pkg.inputs
    a = inp.lattice_scale_lj(0.8442)
    x, cell = inp.fcc_lattice(cells, cells, cells, a)
    v = inp.velocity_create_geom(x, np.ones(len(x)), 1.44, 87287, "lj")
    md = pkg.B200MD("lj", 1)
    md.box([0, 0, 0], [cells * a] * 3)
    md.set_mass(1, 1.0)
    md.pair_style("lj/cut", 2.5).coeff(1, 1, 1.0, 1.0, 2.5)
    md.neighbor.set(0.3)
    md.neighbor.modify(every=20, delay=0, check=0)
    md.create_atoms(x, v, np.ones(len(x), dtype=np.int32), inp.fcc_tags(cell, cells, cells))
    md.run(60)
    g = md.gather()
    md.close()
    return g["x"], cells * a


def max_distinct_per_bank(slots, bank_of):
    """largest number of distinct addresses in one bank"""
    u = np.unique(slots)
    if len(u) == 0:
        return 0
    return int(np.bincount(bank_of(u)).max())


def main():
    cells = int(sys.argv[1]) if len(sys.argv) > 1 else 14
    G = int(sys.argv[2]) if len(sys.argv) > 2 else 6
    x, box = melt(cells)
    cutneigh = 2.8
    nb = int(box / cutneigh)
    bs = box / nb
    b3 = np.minimum((x / bs).astype(int), nb - 1)
    sub = ((x / bs - b3) * 2).astype(int).clip(0, 1)
    binid = (b3[:, 2] * nb + b3[:, 1]) * nb + b3[:, 0]
    order = np.lexsort((np.arange(len(x)), sub[:, 0] + 2 * sub[:, 1] + 4 * sub[:, 2], binid))   # device: bin, then sub-cell order
    xs, bins = x[order], binid[order]
    start = np.searchsorted(bins, np.arange(nb ** 3 + 1))
    G = min(G, nb - 2)
    tot = dict(sectors=0, aos=0, soa=0, xy_z=0, steps=0)
    for bz in range(1, nb - 1):
        for by in range(1, nb - 1):
            bx0 = 1
            rows = []   # (slot range in global sorted order) per r
            for dz in (-1, 0, 1):
                for dy in (-1, 0, 1):
                    first = ((bz + dz) * nb + (by + dy)) * nb + bx0 - 1
                    rows.append((start[first], start[first + G + 2]))
            off = np.concatenate([[0], np.cumsum([e - s for s, e in rows])])
            own = np.arange(start[(bz * nb + by) * nb + bx0], start[(bz * nb + by) * nb + bx0 + G])
            lists = []
            for i in own:
                bxi = bins[i] % nb
                ent = []
                for r, (s, e) in enumerate(rows):
                    k = np.arange(s, e)
                    ok = (np.abs(bins[k] % nb - bxi) <= 1) & (k != i)
                    d = xs[k] - xs[i]
                    ok &= (d * d).sum(1) <= cutneigh * cutneigh
                    ent.append(off[r] + k[ok] - s)
                lists.append(np.concatenate(ent))
            for w0 in range(0, len(own), 32):
                ls = lists[w0:w0 + 32]
                for k in range(max(len(l) for l in ls)):
                    lanes = np.array([l[k] if k < len(l) else -1 for l in ls] + [-1] * (32 - len(ls)))
                    tot["steps"] += 1
                    tot["sectors"] += len(np.unique(lanes[lanes >= 0]))
                    aos = 0
                    for q in range(4):   # LDS.128: quarter-warp phases, a 32-byte record starts at bank 8 * (slot % 4)
                        p = lanes[8 * q:8 * q + 8]
                        aos += max_distinct_per_bank(p[p >= 0], lambda u: u % 4)
                    soa = 0
                    for h in range(2):   # LDS.64: half-warp phases, element `slot` sits in bank pair slot % 16
                        p = lanes[16 * h:16 * h + 16]
                        aos += max_distinct_per_bank(p[p >= 0], lambda u: u % 4)
                        soa += 3 * max_distinct_per_bank(p[p >= 0], lambda u: u % 16)
                    xy_z = soa // 3          # z[]: one LDS.64
                    for q in range(4):        # (x,y)[]: one LDS.128, 16-byte elements -> 8 bank groups
                        p = lanes[8 * q:8 * q + 8]
                        xy_z += max_distinct_per_bank(p[p >= 0], lambda u: u % 8)
                    tot["aos"] += aos
                    tot["soa"] += soa
                    tot["xy_z"] += xy_z
    n = tot["steps"]
    print(f"cells {cells}  bins/dim {nb}  G {G}  warp steps {n}")
    print(f"  L1 gather, distinct sectors per warp step : {tot['sectors'] / n:6.2f}   (ncu on k_lj_lanes: 26 wavefronts)")
    print(f"  shared, 32-byte records (LDS.128 + LDS.64) : {tot['aos'] / n:6.2f} wavefronts per warp step")
    print(f"  shared, x[] y[] z[]      (3 x LDS.64)      : {tot['soa'] / n:6.2f} wavefronts per warp step")
    print(f"  shared, (x,y)[] z[]  (LDS.128 + LDS.64)    : {tot['xy_z'] / n:6.2f} wavefronts per warp step")


if __name__ == "__main__":
    main()
