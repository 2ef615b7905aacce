"""N>1 host logic on CPU: world_size-2 gloo run of the decomposed input generation, and the 26-direction
peer plan of the single-stage halo (b200md_plan_peers is host-only C, no device needed)."""
from __future__ import annotations

import ctypes as C
import importlib

import numpy as np
import pytest

from multirank_util import run_decomposed


def _peers(pkg, grid, periodic, rank):
    L = pkg.lib.load()
    g = (C.c_int * 3)(*grid)
    p = (C.c_int * 3)(*periodic)
    ps, pr = (C.c_int * 26)(), (C.c_int * 26)()
    assert L.b200md_plan_peers(g, p, rank, ps, pr) == 0
    return list(ps), list(pr)


def _dir_vec(d):
    full = d if d < 13 else d + 1
    return (full % 3 - 1, (full // 3) % 3 - 1, full // 9 - 1)


@pytest.mark.parametrize("grid", [(1, 1, 1), (2, 1, 1), (2, 2, 1), (2, 2, 2), (4, 2, 1), (3, 3, 3)])
@pytest.mark.parametrize("periodic", [(1, 1, 1), (1, 0, 1), (0, 0, 0)])
def test_plan_peers_consistent(pkg, grid, periodic):
    """A sends to B in direction d  <=>  B receives direction d from A; geometry matches the rank grid
    (x fastest, like CommBrick's procgrid / myloc, src/comm_brick.cpp:191-260)."""
    n = grid[0] * grid[1] * grid[2]
    plans = [_peers(pkg, grid, periodic, r) for r in range(n)]
    for a in range(n):
        loc = (a % grid[0], (a // grid[0]) % grid[1], a // (grid[0] * grid[1]))
        for d in range(26):
            b = plans[a][0][d]
            dv = _dir_vec(d)
            tgt = [loc[k] + dv[k] for k in range(3)]
            outside = any((t < 0 or t >= grid[k]) and not periodic[k] for k, t in enumerate(tgt))
            if outside:
                assert b == -1
                continue
            tgt = [t % grid[k] for k, t in enumerate(tgt)]
            assert b == (tgt[2] * grid[1] + tgt[1]) * grid[0] + tgt[0]
            assert plans[b][1][d] == a
    assert L_bad_rank(pkg)


def L_bad_rank(pkg):
    L = pkg.lib.load()
    g = (C.c_int * 3)(2, 1, 1)
    buf = (C.c_int * 26)()
    return L.b200md_plan_peers(g, g, 5, buf, buf) == 5  # B200MD_ERR_ARG


def test_gloo_world2_decomposed_inputs_equal_single_rank(pkg):
    """Two gloo ranks each build their half of the lattice (tags, loop-geom velocities with the momentum/temperature
    sums all-reduced): the union equals the single-rank generation bit for bit in x/tag and to round-off in v."""
    import bench

    res = run_decomposed(2, (2, 1, 1), "lj", 6, 0, backend="gloo", port=29631)
    x1, v1, tag1, box1, n1 = bench.make_atoms(pkg, "lj", 6, (1, 1, 1), (0, 0, 0), lambda a: a)  # 6^3 cells: half of the 12x6x6 box
    x = np.concatenate([r["x0"] for r in res])
    v = np.concatenate([r["v0"] for r in res])
    tag = np.concatenate([r["tag0"] for r in res])
    assert int(res[0]["natoms"][0]) == 4 * 12 * 6 * 6 == len(x)
    assert sorted(tag.tolist()) == list(range(1, len(x) + 1))
    # the same global system generated on one rank
    inp = pkg.inputs
    a = inp.lattice_scale_lj(0.8442)
    xs, cell = inp.fcc_lattice(12, 6, 6, a)
    tags = inp.fcc_tags(cell, 12, 6)
    vs = inp.velocity_create_geom(xs, np.ones(len(xs)), 1.44, 87287, "lj")
    o, os_ = np.argsort(tag), np.argsort(tags)
    assert np.array_equal(x[o], xs[os_])
    np.testing.assert_allclose(v[o], vs[os_], rtol=1e-12, atol=1e-14)
    # each rank's atoms lie inside its own sub-domain [lo, hi)
    half = res[0]["box"][0] / 2
    assert (res[0]["x0"][:, 0] < half).all() and (res[1]["x0"][:, 0] >= half).all()
