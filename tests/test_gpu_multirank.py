"""Spatial decomposition over >= 2 B200s (NCCL halo, exchange, collective rebuild vote) against the same
system on one GPU.  Skipped on a single-GPU box."""
from __future__ import annotations

import importlib

import numpy as np
import pytest

from multirank_util import run_decomposed

pytestmark = pytest.mark.gpu


def _ngpu():
    import torch

    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _single(pkg, kind, cells_xyz, steps):
    import bench

    inp = pkg.inputs
    w = bench.WORK[kind]
    a = inp.lattice_scale_lj(0.8442) if kind == "lj" else 3.615
    x, cell = inp.fcc_lattice(*cells_xyz, a)
    tag = inp.fcc_tags(cell, cells_xyz[0], cells_xyz[1])
    v = inp.velocity_create_geom(x, np.full(len(x), w["mass"]), w["T"], w["seed"], w["units"])
    md = bench.make_system(pkg, kind, 0, 1, 0, None, (1, 1, 1))
    md.box([0, 0, 0], [c * a for c in cells_xyz])
    md.create_atoms(x, v, np.ones(len(x), dtype=np.int32), tag)
    md.setup()
    t0 = md.thermo_log[-1]
    nn = int(md.neighbor_list()[0].sum())
    md.run(steps)
    t1 = md.thermo_log[-1]
    g = md.gather()
    st = md.neigh_stats()
    md.close()
    return t0, t1, g, nn, st


@pytest.mark.parametrize("kind,world,grid,cells,steps", [
    ("lj", 2, (2, 1, 1), 8, 60),
    ("eam", 2, (2, 1, 1), 8, 40),
    ("lj", 4, (2, 2, 1), 8, 60),
    ("eam", 8, (2, 2, 2), 6, 40),
])
def test_decomposed_equals_single_gpu(pkg, kind, world, grid, cells, steps):
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    res = run_decomposed(world, grid, kind, cells, steps, backend="nccl", port=29641 + world)
    tot = tuple(cells * g for g in grid)
    t0, t1, g, nn, st = _single(pkg, kind, tot, steps)
    # step 0: same neighbor-pair count, thermo to round-off
    assert sum(int(r["nneigh0"][0]) for r in res) == nn

    np.testing.assert_allclose(res[0]["thermo0"], [t0["pe"], t0["ke"], t0["press"]], rtol=1e-10)
    for r in res[1:]:
        assert np.array_equal(r["thermo0"], res[0]["thermo0"])
    # after the run (rebuilds, migration, halo every step): thermo within north_star's 1e-6, same rebuild count
    np.testing.assert_allclose(res[0]["thermo1"], [t1["pe"], t1["ke"], t1["press"]], rtol=1e-6)
    assert all(int(r["builds"][0]) == st["builds"] for r in res)
    # no atom lost or duplicated; positions agree with the single-GPU trajectory
    tag = np.concatenate([r["tag1"] for r in res])
    x = np.concatenate([r["x1"] for r in res])
    assert sorted(tag.tolist()) == list(range(1, len(g["tag"]) + 1))
    o = np.argsort(tag)
    box = res[0]["box"]
    dx = x[o] - g["x"]
    dx -= box * np.round(dx / box)
    assert np.abs(dx).max() < 1e-6
    assert sum(int(r["nlocal1"][0]) for r in res) == len(tag)


@pytest.mark.parametrize("kind,world,grid,cells,steps", [
    ("lj", 2, (2, 1, 1), 8, 60),
    ("eam", 2, (2, 1, 1), 8, 40),
    ("eam", 4, (2, 2, 1), 8, 40),
    ("lj", 8, (2, 2, 2), 6, 60),
])
def test_halo_put_equals_nccl_halo(pkg, kind, world, grid, cells, steps, monkeypatch):
    """the per-step halo as ONE fused pack-and-put kernel over peer memory (cudaIpc mappings of the peers' landing
    buffers, per-direction flags, wait + unpack on the receiver) against pack + ncclSend/ncclRecv: bit-identical
    trajectories through rebuilds and migration, and the put path really ran"""
    if _ngpu() < world:
        pytest.skip(f"needs {world} GPUs")
    monkeypatch.delenv("B200MD_TEST_HALO_PUT", raising=False)
    a = run_decomposed(world, grid, kind, cells, steps, backend="nccl", port=29671 + world)
    monkeypatch.setenv("B200MD_TEST_HALO_PUT", "400000")
    b = run_decomposed(world, grid, kind, cells, steps, backend="nccl", port=29681 + world)
    for ra, rb in zip(a, b):
        assert int(ra["puts"][0]) == 0 and int(rb["puts"][0]) >= steps // 2
        assert np.array_equal(ra["tag1"], rb["tag1"]) and np.array_equal(ra["x1"], rb["x1"]) and np.array_equal(ra["f1"], rb["f1"])
        assert np.array_equal(ra["thermo1"], rb["thermo1"]) and np.array_equal(ra["builds"], rb["builds"])
