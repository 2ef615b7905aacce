"""Synthetic benchmark inputs (host side, numpy): what `lattice fcc`, `create_atoms`, and
`velocity all create T seed loop geom` produce in bench/in.lj and bench/in.eam of the reference.

Follows src/lattice.cpp:90-94,255-262 (fcc basis, lj-units scale), src/create_atoms.cpp:1301-1370 (loop
k, j, i, basis; tags in creation order on one rank), src/velocity.cpp:329-369 with src/random_park.cpp:23-49
(Park-Miller) and :92-128 (per-atom seed = Jenkins one-at-a-time hash of the seed and coordinate bytes).
"""
from __future__ import annotations

import numpy as np

FCC_BASIS = np.array([[0.0, 0.0, 0.0], [0.5, 0.5, 0.0], [0.5, 0.0, 0.5], [0.0, 0.5, 0.5]])

UNITS = {  # src/update.cpp:145-207
    "lj": dict(boltz=1.0, mvv2e=1.0, ftm2v=1.0, nktv2p=1.0, dt=0.005, skin=0.3),
    "metal": dict(boltz=8.617343e-5, mvv2e=1.0364269e-4, ftm2v=1.0 / 1.0364269e-4, nktv2p=1.6021765e6, dt=0.001, skin=2.0),
    "real": dict(boltz=0.0019872067, mvv2e=48.88821291 * 48.88821291, ftm2v=1.0 / 48.88821291 / 48.88821291,
                 nktv2p=68568.415, dt=1.0, skin=2.0),
}


def lattice_scale_lj(rhostar: float) -> float:
    """lj units: reduced density -> lattice constant, fcc has 4 basis atoms (src/lattice.cpp:255-262)."""
    return float(np.power(4.0 / rhostar, 1.0 / 3.0))


def fcc_lattice(nx: int, ny: int, nz: int, scale: float, lo=(0, 0, 0)):
    """Positions (n,3) and tags (n,) of the fcc cells [lo, lo+n) in creation order k, j, i, basis.
    tag = 1 + 4*((k*ny_total + j)*nx_total + i) + m only holds for lo == 0 on one rank; callers that
    decompose pass global cell indices through `lo` and get global tags back via fcc_tags()."""
    k, j, i, m = np.meshgrid(np.arange(lo[2], lo[2] + nz), np.arange(lo[1], lo[1] + ny), np.arange(lo[0], lo[0] + nx),
                             np.arange(4), indexing="ij")
    x = np.empty((k.size, 3))
    x[:, 0] = ((i + FCC_BASIS[m, 0]) * scale).ravel()
    x[:, 1] = ((j + FCC_BASIS[m, 1]) * scale).ravel()
    x[:, 2] = ((k + FCC_BASIS[m, 2]) * scale).ravel()
    return x, (i.ravel(), j.ravel(), k.ravel(), m.ravel())


def fcc_tags(cell, nx_total: int, ny_total: int):
    i, j, k, m = cell
    return (1 + 4 * ((k.astype(np.int64) * ny_total + j) * nx_total + i) + m).astype(np.int32)


def _park_step(seed):
    k = seed // 127773
    seed = 16807 * (seed - k * 127773) - 2836 * k
    seed = np.where(seed < 0, seed + 2147483647, seed)
    return seed


def velocity_loop_geom_raw(x: np.ndarray, seed: int) -> np.ndarray:
    """Per-atom uniform(-0.5, 0.5) triples exactly as `loop geom` draws them (before 1/sqrt(m), momentum
    removal and rescaling): depends only on the seed and the bit pattern of each atom's coordinates."""
    x = np.ascontiguousarray(x, dtype=np.float64)
    n = x.shape[0]
    h = np.zeros(n, dtype=np.uint32)

    def mix(byte_col):
        nonlocal h
        h = h + byte_col.astype(np.int32).astype(np.uint32)
        h = h + (h << np.uint32(10))
        h = h ^ (h >> np.uint32(6))

    for b in np.array([seed], dtype=np.int32).view(np.int8):
        mix(np.full(n, b, dtype=np.int8))
    xb = x.view(np.int8).reshape(n, 24)
    for c in range(24):
        mix(xb[:, c])
    h = h + (h << np.uint32(3))
    h = h ^ (h >> np.uint32(11))
    h = h + (h << np.uint32(15))
    s = (h & np.uint32(0x7FFFFFF)).astype(np.int64)
    s[s == 0] = 1
    for _ in range(5):
        s = _park_step(s)
    v = np.empty((n, 3))
    for c in range(3):
        s = _park_step(s)
        v[:, c] = (1.0 / 2147483647) * s - 0.5
    return v


def velocity_finish(v: np.ndarray, mass_per_atom: np.ndarray, t_desired: float, units: str, natoms_total=None,
                    allsum=lambda a: a) -> np.ndarray:
    """1/sqrt(m) scaling, zero total momentum, rescale to t_desired (src/velocity.cpp:333-369; dof = 3N-3).
    `allsum` sums an array over ranks for decomposed runs (identity on one rank)."""
    u = UNITS[units]
    m = np.asarray(mass_per_atom, dtype=np.float64)
    v = v / np.sqrt(m)[:, None]
    n_total = natoms_total if natoms_total is not None else v.shape[0]
    tot = allsum(np.concatenate([(v * m[:, None]).sum(axis=0), [m.sum()]]))
    v = v - tot[:3] / tot[3]
    t = allsum(np.array([(m[:, None] * v * v).sum()]))[0]
    dof = 3.0 * n_total - 3.0
    t *= u["mvv2e"] / (dof * u["boltz"])
    return v * np.sqrt(t_desired / t)


def velocity_create_geom(x, mass_per_atom, t_desired, seed, units="lj"):
    return velocity_finish(velocity_loop_geom_raw(x, seed), mass_per_atom, t_desired, units)


def replicate_chain(g: dict, r: int) -> dict:
    """`replicate r r r` (src/replicate.cpp) of the bench/data.chain system as tests/golden/chain_32k.npz holds it
    (in_x, in_v, in_type, in_tag, in_image, in_bonds, in_boxlo, in_boxhi): every replica is the UNWRAPPED molecule set
    (x + image * L) shifted by whole box lengths, tags and bond partners offset by replica * natoms, then wrapped into
    the r-times larger box -- so chains that cross the periodic boundary of the small box stay connected.
    Returns x, v, type, tag, bonds (type, a, b), boxlo, boxhi."""
    lo, hi = np.asarray(g["in_boxlo"], dtype=np.float64), np.asarray(g["in_boxhi"], dtype=np.float64)
    L = hi - lo
    n = len(g["in_tag"])
    xu = np.asarray(g["in_x"], dtype=np.float64) + np.asarray(g["in_image"], dtype=np.float64) * L
    xs, vs, tags, bonds = [], [], [], []
    for k in range(r):
        for j in range(r):
            for i in range(r):
                rep = (k * r + j) * r + i
                xs.append(xu + np.array([i, j, k]) * L)
                vs.append(g["in_v"])
                tags.append(np.asarray(g["in_tag"], dtype=np.int64) + rep * n)
                b = np.array(g["in_bonds"], dtype=np.int64)
                b[:, 1:] += rep * n
                bonds.append(b)
    x = np.concatenate(xs)
    big = L * r
    x = lo + np.mod(x - lo, big)
    x = np.where(x >= lo + big, lo, x)   # mod() can round up to the upper face
    return dict(x=x, v=np.concatenate(vs), type=np.tile(np.asarray(g["in_type"], dtype=np.int32), r ** 3),
                tag=np.concatenate(tags).astype(np.int32), bonds=np.concatenate(bonds).astype(np.int32), boxlo=lo, boxhi=lo + big)


def bond_arrays(tag, bonds):
    """data-file `Bonds` rows (type a1 a2) -> num_bond[n], bond_type[n][bpa], bond_atom[n][bpa] in the local order of
    `tag`, each bond stored on a1 (newton_bond on; Atom::data_bonds src/atom.cpp:1380-1460).  Vectorised: the chain
    benchmark replicates to millions of bonds."""
    tag = np.asarray(tag)
    bonds = np.asarray(bonds)
    order = np.argsort(tag, kind="stable")
    where = order[np.searchsorted(tag[order], bonds[:, 1])]
    num = np.bincount(where, minlength=len(tag)).astype(np.int32)
    bpa = max(1, int(num.max()))
    o = np.argsort(where, kind="stable")
    slot = np.arange(len(o)) - np.repeat(np.cumsum(num) - num, num)
    btype = np.zeros((len(tag), bpa), dtype=np.int32)
    batom = np.zeros((len(tag), bpa), dtype=np.int32)
    btype[where[o], slot] = bonds[o, 0]
    batom[where[o], slot] = bonds[o, 2]
    return num, btype, batom
