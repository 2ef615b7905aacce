"""Kernel and host LOGIC of libb200md's sources, executed on the CPU through the lockstep CUDA emulation shim in
tests/cudaemu/ (TEST INFRASTRUCTURE: the same .cu files, `<<<>>>` rewritten, compiled with g++; fibers for threads,
guard pages around every device buffer).  These tests exist because code written after a round's GPU budget is spent
would otherwise first execute at the round-end GPU run; they check indexing, flag protocols, call order and buffer
sizes against the same goldens and oracle the -m gpu tests use.  They are NOT parity evidence for the device (no FMA
contraction, exact reciprocal): that is what tests marked gpu are for.  The product never loads this library."""
from __future__ import annotations

import ctypes as C
import importlib
import os
import sys
from pathlib import Path

import numpy as np
import pytest

from conftest import CASES, load_golden, make_device_system, thermo_vec

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT / "tests" / "cudaemu"))

F_RTOL = 1e-10
THERMO_RTOL = 1e-6


@pytest.fixture()
def emu(monkeypatch):
    """the package with lib._lib swapped for the emulation build of the same sources (this test process only)"""
    import build as emu_build  # tests/cudaemu/build.py

    path = emu_build.build()
    pkg = importlib.import_module("lammps-fugaku-threadpool_b200")
    L = C.CDLL(str(path))
    for name, (res, args) in pkg.lib.PROTOTYPES.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    monkeypatch.setattr(pkg.lib, "_lib", L)
    monkeypatch.setattr(pkg.lib, "load", lambda rebuild=False: L)
    return pkg


@pytest.mark.parametrize("name", list(CASES))
def test_emu_forces_and_trajectory_vs_reference(emu, name):
    """validates the emulator itself on paths the B200 has already run green: setup forces and the short trajectory
    of every reference-minted golden case"""
    g = load_golden(name)
    md = make_device_system(emu, name, g)
    md.setup()
    a = md.gather()
    assert np.array_equal(a["tag"], g["a_tag"])
    assert np.abs(a["f"] - g["a_f"]).max() <= F_RTOL * np.abs(g["a_f"]).max()
    np.testing.assert_allclose(thermo_vec(md.thermo_log[-1]), g["a_thermo"][:3], rtol=1e-9)
    md.run(int(g["nmore"][0]))
    a = md.gather()
    box = g["boxhi"] - g["boxlo"]
    dx = a["x"] - g["b_x"]
    dx -= box * np.round(dx / box)
    assert np.abs(dx).max() < 1e-7
    np.testing.assert_allclose(thermo_vec(md.thermo_log[-1]), g["b_thermo"][:3], rtol=THERMO_RTOL)
    md.close()


# ---- bench/in.chain device path (bond.cu, BONDED neighbor build, langevin replay) ----
import test_zz_gpu_chain as chain  # noqa: E402  (the GPU tests' bodies, run here against the emulation build)


def test_emu_chain_setup(emu):
    chain.test_chain_setup_forces_vs_reference(emu)


def test_emu_chain_bond_build_variants(emu):
    chain.test_chain_bond_build_variants_agree(emu)


@pytest.mark.parametrize("thermostat", [False, True])
def test_emu_chain_fused_post_force(emu, thermostat):
    chain.test_chain_fused_post_force_equals_separate_kernels(emu, thermostat)


def test_emu_chain_100_steps(emu):
    chain.test_chain_100_steps_vs_reference_and_shipped_log(emu)


def test_emu_chain_error_paths(emu):
    chain.test_chain_error_paths(emu)


# ---- paths of the atomic hot path that were written after the GPU budget was spent (opt-in on the device) ----
import test_gpu_parity as gp  # noqa: E402


@pytest.mark.parametrize("pipeline", [False, True])
@pytest.mark.parametrize("name,steps", [("lj2_864", 25), ("eam_500", 30)])
def test_emu_host_buffer_step(emu, name, steps, pipeline):
    """b200md_step_host, serial form and the three-stream chunk pipeline (the default; the emulation runs streams in
    issue order, so this checks its indexing and chunking, not its event protocol)"""
    gp.test_host_buffer_step_equals_device_loop(emu, name, steps, pipeline)


@pytest.mark.parametrize("kind,steps", [("lj", 45), ("eam", 25)])
def test_emu_host_buffer_pipeline_chunked(emu, oracle, kind, steps):
    gp.test_host_buffer_pipeline_chunked(emu, oracle, kind, steps)


@pytest.mark.parametrize("name,steps", [("lj_500", 45), ("eam_500", 30)])
def test_emu_overlapped_halo(emu, name, steps):
    """overlap mode: interior / boundary tile split of the force kernels (tile_ghost flags from the build)"""
    gp.test_overlapped_halo_equals_plain(emu, name, steps)


@pytest.mark.parametrize("name", list(CASES))
def test_emu_neighbor_list_bit_exact(emu, name):
    gp.test_neighbor_list_bit_exact(emu, name)
    for gb in (2, 5):                                # several bins of an x-row per warp (sparse-bin form)
        gp.test_neighbor_list_bit_exact(emu, name, tune={"build_gb": gb})
    gp.test_neighbor_list_bit_exact(emu, name, tune={"build_fp64": 1})   # the all-FP64 fallback kernel
    gp.test_neighbor_list_bit_exact(emu, name, tune={"build_seg_cap": 8})  # staging area too small: grow-and-redo path


@pytest.mark.parametrize("kind,steps", [("lj", 45), ("eam", 25)])
def test_emu_bins_per_warp_same_list_order(emu, oracle, kind, steps):
    """k_build_seg with segments of 1, 3 and 7 bins per block (the box has 4 bins per x-row: 3 leaves a remainder
    segment, 7 is clamped to the whole row and needs several passes of the 128 threads) writes the same lists in the
    same entry order: trajectories are bit-identical"""
    out = []
    for gb in (1, 3, 7):
        md, _ = gp._bench_system(emu, oracle, kind, 8)
        md.tune("build_gb", gb)
        md.run(steps)
        assert emu.lib.load().b200md_build_bins_per_block(md.ctx.h) == min(int(gb), 4)
        out.append((md.gather(), md.thermo_log[-1]["pe"], md.neigh_stats()))
        md.close()
    for g, pe, st in out[1:]:
        assert st == out[0][2] and st["builds"] >= 2 and pe == out[0][1]
        assert np.array_equal(g["x"], out[0][0]["x"]) and np.array_equal(g["f"], out[0][0]["f"])


@pytest.mark.parametrize("ntypes", [1, 3])
def test_emu_dilute_gas_and_special_bonds_direct_sums(emu, ntypes):
    """oracle-independent checks: a dilute gas (empty and ragged bins, vector width 2, one and three types) and, once,
    the special-bond path through the C ABI, each against a direct O(N^2) sum"""
    gp.test_dilute_gas_vs_direct_sum(emu, ntypes)
    if ntypes == 1:
        gp.test_host_ghost_special_bonds_vs_direct_sum(emu)


@pytest.mark.parametrize("name,steps", [("lj_500", 45), ("eam_500", 30)])
@pytest.mark.parametrize("overlap", [0, 1])
def test_emu_fused_nve_equals_separate_kernels(emu, name, steps, overlap):
    gp.test_fused_nve_equals_separate_kernels(emu, name, steps, overlap)


def test_emu_piecewise_api_and_error_paths(emu):
    gp.test_download_in_the_callers_order(emu)
    gp.test_piecewise_api_equals_device_loop(emu)
    gp.test_error_paths(emu)
    gp.test_run_to_run_bitwise_reproducible(emu)


@pytest.mark.parametrize("kind,n,steps", [("lj", 12, 45), ("eam", 12, 25)])
def test_emu_overlap_with_interior_tiles(emu, oracle, kind, n, steps):
    """a box wide enough (12^3 cells) that many 32-atom tiles have no ghost in their lists, so the interior part of
    the split really runs under the halo: overlap == plain, bit for bit, through rebuilds"""
    out = []
    for overlap in (False, True):
        md, _ = gp._bench_system(emu, oracle, kind, n)
        md.tune("overlap", int(overlap))
        md.run(steps)
        out.append((md.gather(), md.thermo_log[-1]["pe"], md.neigh_stats()["builds"]))
        md.close()
    (a, ea, ba), (b, eb, bb) = out
    assert ba == bb and ba >= 2
    assert np.array_equal(a["x"], b["x"]) and np.array_equal(a["v"], b["v"]) and np.array_equal(a["f"], b["f"])
    assert ea == eb


@pytest.mark.parametrize("kind,n,steps", [("lj", 10, 100), ("eam", 10, 100)])
def test_emu_bench_inputs_vs_oracle(emu, oracle, kind, n, steps):
    gp.test_bench_inputs_vs_oracle(emu, oracle, kind, n, steps)


# ---- the LAMMPS-side glue (lmp_b200 = unmodified reference + b200 style classes) on the emulation build ----
import test_lammps_host as lh  # noqa: E402


@pytest.fixture()
def emu_lmp(monkeypatch):
    """lmp_b200 finds libb200md.so through RUNPATH, which LD_LIBRARY_PATH precedes: point it at a directory where
    that name is the emulation build"""
    import build as emu_build

    path = emu_build.build()
    lh.build_host()
    if not lh.LMP_REF.exists():
        pytest.skip("lmp_ref not built")
    d = path.parent / "as_libb200md"
    d.mkdir(exist_ok=True)
    link = d / "libb200md.so"
    if not link.is_symlink():
        link.symlink_to(path)
    monkeypatch.setenv("LD_LIBRARY_PATH", f"{d}:{os.environ.get('LD_LIBRARY_PATH', '')}")


@pytest.mark.parametrize("kind,n", [("lj", 10), ("eam", 10), ("lj", 7)])
@pytest.mark.parametrize("loop", ["verlet/b200", "verlet"])
def test_emu_lammps_thermo_matches_reference(emu_lmp, kind, n, loop, tmp_path):
    lh.test_thermo_matches_reference(kind, n, loop, tmp_path)


@pytest.mark.parametrize("kind,n,steps", [("lj", 6, 25), ("eam", 6, 12)])
def test_emu_lammps_device_list_through_library_api(emu_lmp, kind, n, steps, tmp_path):
    lh.check_device_list_through_library_api(tmp_path, kind, n, steps)


@pytest.mark.parametrize("name,units", [("unittest_atomic_pair_eam", "metal"), ("unittest_atomic_pair_eam_real", "real")])
def test_emu_lammps_unittest_eam_yaml(emu_lmp, name, units, tmp_path):
    """host-ghost mode on a box thinner than the cutoff (several periodic image layers made by LAMMPS), two
    elements: the reference's own known-answer file through pair eam/b200"""
    lh.test_reference_unittest_eam_yaml_through_lammps(name, units, tmp_path)


def test_emu_lammps_unittest_mol_pair_lj_cut_yaml(emu_lmp, tmp_path):
    """molecular system under the stock loop: special-bond lists from the host applied in the device list build, factor_lj
    in the multi-type lj kernel -- the reference's mol-pair-lj_cut.yaml"""
    lh.test_reference_unittest_mol_pair_lj_cut_yaml_through_lammps(tmp_path)


def test_emu_lammps_thin_box_refused(emu_lmp, tmp_path):
    lh.test_thin_box_is_refused_by_the_resident_loop(tmp_path)


def test_emu_lammps_chain(emu_lmp, tmp_path):
    """bench/in.chain with -sf b200 end to end through the glue: VerletB200::setup on the device, bond fene/b200,
    fix langevin/b200, energies published to the reference's computes"""
    lh.test_chain_thermo_matches_reference(tmp_path)


# ---- several ranks: threads of this process, one emulated device each, NCCL stand-in (tests/cudaemu/fake_nccl.cpp) ----
import threading  # noqa: E402


def emu_build_dir():
    import build as emu_build

    return emu_build.OUT


class _AllSum:
    def __init__(self, n):
        self.n, self.bar, self.parts, self.lock = n, threading.Barrier(n, timeout=300), {}, threading.Lock()

    def __call__(self, rank, a):
        a = np.ascontiguousarray(a, dtype=np.float64)
        with self.lock:
            self.parts[rank] = a
        self.bar.wait()
        tot = sum(self.parts[r] for r in range(self.n))
        self.bar.wait()
        return tot


class _AllGatherBytes:
    """the host-side all-gather of the halo-put handles, ranks being threads"""
    def __init__(self, n):
        self.n, self.bar, self.parts, self.lock = n, threading.Barrier(n, timeout=300), {}, threading.Lock()

    def __call__(self, rank, blob):
        with self.lock:
            self.parts[rank] = blob
        self.bar.wait()
        out = [self.parts[r] for r in range(self.n)]
        self.bar.wait()
        return out


def run_decomposed_emu(pkg, world, grid, kind, cells, steps, host_steps=0, configure=None, tune=None):
    """tests/multirank_util._worker with threads for ranks; optionally `host_steps` more steps through
    b200md_step_host (the bench's e2e leg: atoms migrate between ranks at its rebuilds)"""
    import bench

    allsum, res, errs = _AllSum(world), [None] * world, []
    nccl_id = pkg.lib.nccl_unique_id()

    def worker(rank):
        try:
            loc = (rank % grid[0], (rank // grid[0]) % grid[1], rank // (grid[0] * grid[1]))
            x, v, tag, box, natoms = bench.make_atoms(pkg, kind, cells, grid, loc, lambda a: allsum(rank, a))
            md = bench.make_system(pkg, kind, rank, world, rank, nccl_id, grid)
            md.box([0, 0, 0], box)
            md.create_atoms(x, v, np.ones(len(x), dtype=np.int32), tag, natoms_total=natoms)
            if configure:
                configure(md)
            for k, val in (tune or {}).items():
                md.tune(k, val)
            md.setup()
            t0 = md.thermo_log[-1]
            numneigh, _ = md.neighbor_list()
            out = dict(box=np.array(box), nneigh0=int(numneigh.sum()), thermo0=np.array([t0["pe"], t0["ke"], t0["press"]]),
                       nghost0=md.nghost)
            md.run(steps)
            t1 = md.thermo_log[-1]
            g = md.gather()
            st = md.neigh_stats()
            out.update(thermo1=np.array([t1["pe"], t1["ke"], t1["press"]]), x1=g["x"], tag1=g["tag"], builds=st["builds"],
                       nlocal1=md.nlocal, temp0=t0["temp"], temp1=t1["temp"], puts=int(md.ctx.L.b200md_halo_put_count(md.ctx.h)))
            if host_steps:
                n = md.nlocal
                cap = n + n // 8 + 64
                hx, hv, hf = np.zeros((cap, 3)), np.zeros((cap, 3)), np.zeros((cap, 3))
                htag = np.zeros(cap, dtype=np.int32)
                vp = lambda a: a.ctypes.data_as(C.c_void_p)
                md.ctx.call("b200md_download_atoms", vp(hx), vp(hv), None, None, None, None)
                nout, seen = C.c_int(0), set()
                for _ in range(host_steps):
                    md.ctx.call("b200md_step_host", md.dt, 0, 0, cap, vp(hx), vp(hv), vp(hf), vp(htag), C.byref(nout), None, None)
                    seen.add(nout.value)
                out.update(hx=hx[:nout.value].copy(), htag=htag[:nout.value].copy(), nlocal_seen=sorted(seen),
                           builds2=md.neigh_stats()["builds"])
            md.close()
            res[rank] = out
        except BaseException as e:  # noqa: BLE001
            errs.append((rank, e))
            print(f"emulated rank {rank} failed: {e!r}", file=sys.stderr, flush=True)  # the others will time out in a collective
            allsum.bar.abort()
            C.CDLL(str(Path(emu_build_dir()) / "libnccl_emu.so")).fake_nccl_break_all()
            raise

    th = [threading.Thread(target=worker, args=(r,)) for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs, errs
    return res


@pytest.mark.parametrize("kind,world,grid,cells,steps", [
    ("lj", 2, (2, 1, 1), 6, 60),
    ("eam", 2, (2, 1, 1), 6, 40),
    ("lj", 4, (2, 2, 1), 6, 60),
    ("eam", 8, (2, 2, 2), 5, 40),
])
def test_emu_decomposed_equals_single(emu, kind, world, grid, cells, steps):
    """the body of tests/test_gpu_multirank.py::test_decomposed_equals_single_gpu on emulated ranks"""
    import test_gpu_multirank as mr

    res = run_decomposed_emu(emu, world, grid, kind, cells, steps)
    tot = tuple(cells * g for g in grid)
    t0, t1, g, nn, st = mr._single(emu, kind, tot, steps)
    assert sum(r["nneigh0"] for r in res) == nn
    np.testing.assert_allclose(res[0]["thermo0"], [t0["pe"], t0["ke"], t0["press"]], rtol=1e-10)
    np.testing.assert_allclose(res[0]["thermo1"], [t1["pe"], t1["ke"], t1["press"]], rtol=1e-6)
    assert all(r["builds"] == st["builds"] for r in res)
    tag = np.concatenate([r["tag1"] for r in res])
    x = np.concatenate([r["x1"] for r in res])
    assert sorted(tag.tolist()) == list(range(1, len(g["tag"]) + 1))
    o = np.argsort(tag)
    dx = x[o] - g["x"]
    dx -= res[0]["box"] * np.round(dx / res[0]["box"])
    assert np.abs(dx).max() < 1e-6


@pytest.mark.parametrize("kind,world,grid,cells,steps,host_steps", [
    ("eam", 4, (2, 2, 1), 5, 10, 40),
    ("eam", 8, (2, 2, 2), 5, 10, 40),
    ("lj", 8, (2, 2, 2), 5, 10, 50),
])
def test_emu_host_buffer_step_with_migration(emu, kind, world, grid, cells, steps, host_steps):
    """bench.py's e2e leg on several ranks: b200md_step_host through rebuilds at which atoms change rank (the leg that
    failed on 4 and 8 B200s in round 1 with stale atom counts).  The owned count must really change on some rank,
    no atom may be lost or doubled, and the trajectory must follow the single-rank run of the same total length."""
    import test_gpu_multirank as mr

    res = run_decomposed_emu(emu, world, grid, kind, cells, steps, host_steps=host_steps)
    tot = tuple(cells * g for g in grid)
    _, _, g, _, st = mr._single(emu, kind, tot, steps + host_steps)
    tag = np.concatenate([r["htag"] for r in res])
    x = np.concatenate([r["hx"] for r in res])
    assert sorted(tag.tolist()) == list(range(1, len(g["tag"]) + 1))
    assert any(len(r["nlocal_seen"]) > 1 for r in res), "no migration happened: the case does not test what it claims"
    assert all(r["builds2"] == st["builds"] for r in res)
    o = np.argsort(tag)
    dx = x[o] - g["x"]
    dx -= res[0]["box"] * np.round(dx / res[0]["box"])
    assert np.abs(dx).max() < 1e-6


@pytest.mark.parametrize("kind,world,grid,cells,steps", [("lj", 4, (2, 2, 1), 6, 45), ("eam", 8, (2, 2, 2), 5, 30)])
def test_emu_decomposed_overlap_equals_plain(emu, kind, world, grid, cells, steps):
    """overlap mode on several ranks (its default there; the halo it hides is an NCCL exchange) against the serialized
    halo: same trajectory, bit for bit"""
    a = run_decomposed_emu(emu, world, grid, kind, cells, steps, tune={"overlap": 0})
    b = run_decomposed_emu(emu, world, grid, kind, cells, steps, tune={"overlap": 1})
    for ra, rb in zip(a, b):
        assert np.array_equal(ra["tag1"], rb["tag1"]) and np.array_equal(ra["x1"], rb["x1"])
        assert np.array_equal(ra["thermo1"], rb["thermo1"]) and ra["builds"] == rb["builds"]


@pytest.mark.parametrize("kind,world,grid,cells,steps,overlap", [("lj", 4, (2, 2, 1), 6, 45, 1), ("eam", 8, (2, 2, 2), 5, 30, 1),
                                                                   ("eam", 2, (2, 1, 1), 6, 30, 0)])
def test_emu_halo_put_equals_nccl_halo(emu, kind, world, grid, cells, steps, overlap):
    """the per-step halo as the fused pack-and-put kernel over peer memory (flags, two landing buffers by parity, wait +
    unpack on the receiver) against the NCCL send/recv halo: same trajectory, bit for bit, through rebuilds and
    migration; a landing buffer too small for the ghosts makes every rank fall back to NCCL together"""
    a = run_decomposed_emu(emu, world, grid, kind, cells, steps, tune={"overlap": overlap})
    for cap in (200000, 16):
        gather = _AllGatherBytes(world)
        b = run_decomposed_emu(emu, world, grid, kind, cells, steps, tune={"overlap": overlap},
                               configure=lambda md: md.halo_put(cap, lambda blob: gather(md.rank, blob)))
        for ra, rb in zip(a, b):
            assert np.array_equal(ra["tag1"], rb["tag1"]) and np.array_equal(ra["x1"], rb["x1"])
            assert np.array_equal(ra["thermo1"], rb["thermo1"]) and ra["builds"] == rb["builds"]
            assert ra["puts"] == 0 and (rb["puts"] >= steps // 2 if cap > 16 else rb["puts"] == 0)
    if kind == "eam" and world == 8:
        # landing buffers that hold the ghosts of some list epochs and not of others (the ghost count moves by a few
        # atoms from rebuild to rebuild): the ranks switch between the put halo and NCCL together, parity carries on
        cap = int(np.median([r["nghost0"] for r in a]))
        gather = _AllGatherBytes(world)
        b = run_decomposed_emu(emu, world, grid, kind, cells, steps, tune={"overlap": overlap},
                               configure=lambda md: md.halo_put(cap, lambda blob: gather(md.rank, blob)))
        for ra, rb in zip(a, b):
            assert np.array_equal(ra["tag1"], rb["tag1"]) and np.array_equal(ra["x1"], rb["x1"])
            assert np.array_equal(ra["thermo1"], rb["thermo1"]) and ra["builds"] == rb["builds"]
        assert len({r["puts"] for r in b}) == 1   # every rank served the same halos by put
        print("put halos with a tight landing buffer:", b[0]["puts"], "of", 2 * steps)


@pytest.mark.parametrize("kind,world,grid", [("lj", 2, (2, 1, 1)), ("eam", 4, (2, 2, 1))])
def test_emu_bench_decomposed_parity_leg(emu, kind, world, grid):
    """bench.py's N > 1 parity leg (a small decomposed system against the unmodified reference on the host: thermo at
    step 0 and after 45 steps within 1e-6, same rebuild count, every tag owned exactly once), ranks as threads"""
    import bench

    if not __import__("oracle").ref_available():
        pytest.skip("oracle/_ref not built")
    allsum, res, errs = _AllSum(world), [None] * world, []
    nccl_id = emu.lib.nccl_unique_id()

    def worker(rank):
        try:
            loc = (rank % grid[0], (rank // grid[0]) % grid[1], rank // (grid[0] * grid[1]))
            res[rank] = bench.run_decomposed_parity(None, None, emu, kind, rank, world, rank, lambda: nccl_id, grid, loc,
                                                    lambda a: allsum(rank, a), steps=45, cells=6)
        except Exception as e:  # noqa: BLE001
            errs.append(e)
            allsum.bar.abort()

    th = [threading.Thread(target=worker, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join() for t in th]
    assert not errs, errs
    blk = res[0]
    assert all(r is None for r in res[1:])
    assert blk["ok"], blk
    assert blk["tags"]["owned_exactly_once"] and blk["builds"]["equal"] and blk["builds"]["b200"] >= 2
    assert max(blk["pe"], blk["ke"], blk["press"]) <= 1e-6


def test_emu_langevin_counter_mode_is_rank_count_independent(emu):
    """fix langevin with the counter-based device generator (Philox keyed by seed, counter = tag, step) on an LJ melt:
    1 rank and 4 ranks follow the same trajectory (the numbers depend on nothing but seed, tag and step), and the
    thermostat pulls the temperature from 1.44 toward its 0.3 target"""
    kind, cells, steps = "lj", 6, 60

    def thermostat(md):
        md.fix("langevin", 0.3, 0.3, 0.1, 48279, rng="counter")

    one = run_decomposed_emu(emu, 1, (1, 1, 1), kind, cells * 2, steps, configure=thermostat)[0]
    plain = run_decomposed_emu(emu, 1, (1, 1, 1), kind, cells * 2, steps)[0]
    assert one["temp1"] < 0.75 * plain["temp1"]          # visibly thermostatted toward 0.3
    res = run_decomposed_emu(emu, 8, (2, 2, 2), kind, cells, steps, configure=thermostat)
    np.testing.assert_allclose(res[0]["thermo1"], one["thermo1"], rtol=1e-6)
    tag = np.concatenate([r["tag1"] for r in res])
    x = np.concatenate([r["x1"] for r in res])
    o = np.argsort(tag)
    dx = x[o] - one["x1"]
    dx -= one["box"] * np.round(dx / one["box"])
    assert np.abs(dx).max() < 1e-6


def test_emu_langevin_counter_mode_runs_past_a_sort_step_with_migration(emu):
    """counter-mode thermostat on 2 ranks with `atom_modify sort 10`: atoms migrate between the ranks before the first
    rebuild at a sort step.  The RanMars replay order (sized for the initial owned atoms) must not be consulted in
    counter mode -- it would be stale after the first migration (round-1 advisor finding)."""
    def thermostat(md):
        md.fix("langevin", 1.44, 1.44, 0.5, 48279, rng="counter")
        md.atom_modify_sort(10)

    res = run_decomposed_emu(emu, 2, (2, 1, 1), "lj", 6, 45, configure=thermostat, host_steps=0)
    assert sum(r["nlocal1"] for r in res) == 2 * 4 * 6 ** 3
    assert res[0]["builds"] >= 2 and len({r["nlocal1"] for r in res}) > 1 or res[0]["nlocal1"] != 4 * 6 ** 3


def test_emu_tune_and_special_bonds_errors(emu):
    """b200md_tune refuses unknown keys; a bonded system whose pair style was given special_lj other than the fene
    weights is refused at setup (the bonded build implements exactly lj 0 1 1)"""
    g = load_golden("chain_32k")
    md = chain.make_chain_device(emu, g, langevin=False)
    with pytest.raises(emu.lib.B200MDError) as e:
        md.tune("no_such_knob", 1)
    assert "unknown key" in str(e.value)
    md.pair.special_lj[:] = [1.0, 0.0, 0.0, 0.0]   # the LAMMPS default, not `special_bonds fene`
    md._initialized = False
    with pytest.raises(emu.lib.B200MDError) as e:
        md.setup()
    assert "special_bonds fene" in str(e.value)
    md.close()


def test_emu_chain_with_counter_thermostat(emu):
    """in.chain with the device generator instead of the RanMars replay: runs through the bonded rebuilds, keeps the
    melt at its temperature (T = 1 +- a few %), reproducible run to run"""
    g = load_golden("chain_32k")
    out = []
    for _ in range(2):
        md = chain.make_chain_device(emu, g, langevin=False)
        md.fix("langevin", 1.0, 1.0, 10.0, 904297, rng="counter")
        md.run(8)
        out.append((md.thermo_log[-1], md.gather()))
        md.close()
    (ta, ga), (tb, gb) = out
    assert abs(ta["temp"] - 1.0) < 0.05 and ta["ebond"] > 0
    assert np.array_equal(ga["x"], gb["x"]) and ta["pe"] == tb["pe"]


# ---- eam/alloy and eam/fs (the library's setfl / fs readers + the multi-type EAM kernels) ----
import test_zz_gpu_eam_alloy as alloy  # noqa: E402


@pytest.mark.parametrize("kind,units", [("alloy", "metal"), ("alloy", "real"), ("fs", "metal"), ("fs", "real")])
def test_emu_eam_alloy_fs_vs_oracle(emu, oracle, kind, units, tmp_path):
    alloy.test_eam_alloy_fs_vs_oracle(emu, oracle, kind, units, tmp_path)


def test_emu_eam_alloy_error_paths(emu, tmp_path):
    alloy.test_eam_alloy_error_paths(emu, tmp_path)


@pytest.mark.parametrize("name,units,style", [("unittest_atomic_pair_eam_alloy", "metal", "eam/alloy"),
                                              ("unittest_atomic_pair_eam_alloy_real", "real", "eam/alloy"),
                                              ("unittest_atomic_pair_eam_fs", "metal", "eam/fs"),
                                              ("unittest_atomic_pair_eam_fs_real", "real", "eam/fs")])
def test_emu_lammps_unittest_eam_alloy_fs_yaml(emu_lmp, name, units, style, tmp_path):
    lh.test_reference_unittest_eam_alloy_fs_yaml_through_lammps(name, units, style, tmp_path)


def test_emu_chain_decomposed_equals_single(emu):
    """in.chain on 8 emulated ranks (every rank holds the complete bond table by tag, its own atoms, the counter-based
    thermostat) == the same run on one rank: bonds whose partner lives on another rank resolve through ghost tags,
    tables need no update when beads migrate, bond energy / virial are summed over the ranks"""
    g = load_golden("chain_32k")
    steps, world, grid = 30, 8, (2, 2, 2)
    lo, hi = g["in_boxlo"], g["in_boxhi"]
    num, btype, batom = chain.bond_arrays(g["in_tag"], g["in_bonds"])

    def build(rank, nranks, nccl_id, procgrid, sel):
        md = emu.B200MD(units="lj", ntypes=1, device=rank, nranks=nranks, rank=rank, nccl_id=nccl_id, procgrid=procgrid)
        md.box(lo, hi)
        md.set_mass(1, 1.0)
        p = md.pair_style("lj/cut", 1.12)
        p.coeff(1, 1, 1.0, 1.0, 1.12)
        p.modify(shift=True)
        md.neighbor.set(0.4, "bin")
        md.neighbor.modify(every=1, delay=1, check=True)
        md.timestep(0.012)
        md.create_atoms(g["in_x"][sel], g["in_v"][sel], g["in_type"][sel], g["in_tag"][sel], natoms_total=len(g["in_tag"]))
        md.special_bonds("fene")
        md.bonds(g["in_tag"], num, btype, batom)          # the COMPLETE topology on every rank
        md.bond_style("fene", 1).coeff(1, 30.0, 1.5, 1.0, 1.0)
        md.fix("langevin", 1.0, 1.0, 10.0, 904297, rng="counter")
        return md

    one = build(0, 1, None, (1, 1, 1), slice(None))
    one.run(steps)
    t1, g1 = one.thermo_log[-1], one.gather()
    one.close()

    frac = (g["in_x"] - lo) / (hi - lo)
    cell = np.minimum((frac * np.array(grid)).astype(int), np.array(grid) - 1)
    owner = cell[:, 0] + grid[0] * (cell[:, 1] + grid[1] * cell[:, 2])
    nccl_id, res, errs = emu.lib.nccl_unique_id(), [None] * world, []

    def worker(rank):
        try:
            md = build(rank, world, nccl_id, grid, owner == rank)
            md.run(steps)
            res[rank] = (md.thermo_log[-1], md.gather(), md.neigh_stats()["builds"])
            md.close()
        except BaseException as e:  # noqa: BLE001
            errs.append((rank, e))
            print(f"emulated rank {rank} failed: {e!r}", file=sys.stderr, flush=True)
            C.CDLL(str(Path(emu_build_dir()) / "libnccl_emu.so")).fake_nccl_break_all()
            raise

    th = [threading.Thread(target=worker, args=(r,)) for r in range(world)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs, errs
    t8 = res[0][0]
    for k in ("evdwl", "ebond", "ke", "temp", "press"):
        assert abs(t8[k] - t1[k]) <= 1e-6 * abs(t1[k]), k
    tag = np.concatenate([r[1]["tag"] for r in res])
    x = np.concatenate([r[1]["x"] for r in res])
    assert sorted(tag.tolist()) == list(range(1, len(g["in_tag"]) + 1))
    o = np.argsort(tag)
    dx = x[o] - g1["x"]
    dx -= (hi - lo) * np.round(dx / (hi - lo))
    assert np.abs(dx).max() < 1e-6


def test_emu_replicated_chain_and_bench_leg(emu, monkeypatch):
    """(1) inputs.replicate_chain(2): 256 000 beads whose step-0 energies are exactly 8 x the reference's for the 32 000
    (a periodic system replicated is the same physics), same pressure, no bond longer than R0; (2) bench.py's opt-in
    chain leg wired end to end (the CUDA-event timer replaced by a plain run)"""
    import types

    import bench

    g = load_golden("chain_32k")
    R = emu.inputs.replicate_chain(g, 2)
    md = emu.B200MD(units="lj", ntypes=1)
    md.box(R["boxlo"], R["boxhi"])
    md.set_mass(1, 1.0)
    p = md.pair_style("lj/cut", 1.12)
    p.coeff(1, 1, 1.0, 1.0, 1.12)
    p.modify(shift=True)
    md.neighbor.set(0.4, "bin")
    md.neighbor.modify(every=1, delay=1, check=True)
    md.timestep(0.012)
    md.create_atoms(R["x"], R["v"], R["type"], R["tag"])
    md.special_bonds("fene")
    md.bonds(R["tag"], *emu.inputs.bond_arrays(R["tag"], R["bonds"]))
    md.bond_style("fene", 1).coeff(1, 30.0, 1.5, 1.0, 1.0)
    md.setup()
    t = md.thermo_log[-1]
    evdwl, ebond, ke, temp, press = g["a_thermo"]
    np.testing.assert_allclose([t["evdwl"], t["ebond"], t["ke"]], [8 * evdwl, 8 * ebond, 8 * ke], rtol=1e-10)
    np.testing.assert_allclose(t["press"], press, rtol=1e-10)
    md.close()

    def plain_timer(torch, dist, md, steps, warmup, device, nranks):
        l0 = md.timers()["launches"]
        md.ctx.call("b200md_run", warmup + steps, md.dt, 0)
        return 1.0, md.timers()["launches"] - l0

    monkeypatch.setattr(bench, "time_resident", plain_timer)
    out = bench.run_chain(None, None, emu, 1, types.SimpleNamespace(steps=3, warmup=1), 0)
    assert out["natoms"] == 32000 and out["neigh"]["builds"] >= 1 and out["launches"] > 0
    assert abs(out["thermo1"]["temp"] - 1.0) < 0.05 and out["kernel_ms"]["neigh"][1] >= 0


@pytest.mark.parametrize("kind", ["lj", "eam"])
@pytest.mark.parametrize("loop", ["verlet/b200", "verlet"])
def test_emu_lammps_peratom(emu_lmp, kind, loop, tmp_path):
    lh.test_peratom_energy_and_stress_match_reference(kind, loop, tmp_path)


import test_zz_gpu_peratom as pa  # noqa: E402


@pytest.mark.parametrize("name", list(CASES))
def test_emu_peratom_sums(emu, name):
    pa.test_peratom_sums_are_the_global_tallies(emu, name)


def test_emu_peratom_needs_a_list(emu):
    pa.test_peratom_needs_a_list(emu)


def test_emu_results_do_not_depend_on_block_or_thread_order():
    """CUDA promises no execution order between blocks, nor between the threads of a block outside its barriers: a
    subset of the cases (reference goldens, bit-exact lists, the bonded build, host-buffer steps) once more with
    the emulator running blocks and threads last-to-first (CUDAEMU_ORDER=reverse, read once per process)"""
    import subprocess

    env = dict(os.environ, CUDAEMU_ORDER="reverse")
    r = subprocess.run([sys.executable, "-m", "pytest", str(Path(__file__)), "-x", "-q", "-p", "no:cacheprovider", "-k",
                        "forces_and_trajectory or neighbor_list_bit_exact or chain_setup or emu_host_buffer_step or emu_peratom_sums "
                        "or bins_per_warp or pipeline_chunked"],
                       cwd=ROOT, env=env, capture_output=True, text=True)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert " passed" in r.stdout and "failed" not in r.stdout


def test_emu_lammps_restart_and_data_round_trip(emu_lmp, tmp_path):
    lh.test_restart_and_data_round_trip_under_suffix_b200(tmp_path)


def test_emu_graft_entry_smoke(emu):
    """__graft_entry__.smoke() as the driver runs it at round end, on the emulation build"""
    import __graft_entry__ as ge

    ge.smoke()


class _FakeTorch:
    """the slice of torch.cuda bench.py uses for timing and pinned buffers, on the host: lets bench.main() run end to
    end against the emulation build so that its control flow and its JSON line are exercised before the GPU run"""

    float64 = np.float64
    uint8 = np.uint8

    class _T:
        def __init__(self, a):
            self.a = np.ascontiguousarray(a)

        def pin_memory(self):
            return self

        def data_ptr(self):
            return self.a.ctypes.data

        def item(self):
            return self.a.reshape(-1)[0].item()

        def __getitem__(self, k):
            return _FakeTorch._T(self.a[k:k + 1]) if isinstance(k, int) else _FakeTorch._T(self.a[k])

        def clone(self):
            return _FakeTorch._T(self.a.copy())

    class _Event:
        def __init__(self, enable_timing=False):
            self.t = 0.0

        def record(self):
            import time
            self.t = time.perf_counter()

        def elapsed_time(self, other):
            return (other.t - self.t) * 1e3

    class _Ctx:
        def __enter__(self):
            return self

        def __exit__(self, *a):
            return False

    class cuda:  # noqa: N801
        Event = None

        @staticmethod
        def is_available():
            return True

        @staticmethod
        def set_device(i):
            pass

        @staticmethod
        def synchronize(device=None):
            pass

        @staticmethod
        def ExternalStream(ptr, device=None):  # noqa: N802
            return object()

        @staticmethod
        def stream(s):
            return _FakeTorch._Ctx()

    @staticmethod
    def empty(shape, dtype=None):
        return _FakeTorch._T(np.zeros(shape, dtype=dtype or np.float64))

    @staticmethod
    def tensor(v, dtype=None, device=None):
        return _FakeTorch._T(np.array(v, dtype=dtype or np.float64))


_FakeTorch.cuda.Event = _FakeTorch._Event


def test_emu_bench_main_prints_the_contract_line(emu, monkeypatch, capsys, tmp_path):
    """bench.py, b200 arm, N = 1, reduced sizes: both workloads' resident phases, the e2e phase under its watchdog, the
    cpu_baseline leg (the oracle/_ref reference run), one JSON line with every key the contract names"""
    import json
    import types

    import bench

    fake_dist = types.ModuleType("torch.distributed")
    monkeypatch.setattr(_FakeTorch, "distributed", fake_dist, raising=False)
    monkeypatch.setitem(sys.modules, "torch", _FakeTorch)
    monkeypatch.setitem(sys.modules, "torch.distributed", fake_dist)
    monkeypatch.setattr(sys, "argv", ["bench.py", "--steps", "6", "--warmup", "3", "--cells", "6", "--eam-cells", "6", "--chain", "1"])
    monkeypatch.chdir(tmp_path)
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK"):
        monkeypatch.delenv(k, raising=False)
    bench.main()
    lines = [l for l in capsys.readouterr().out.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    line = json.loads(lines[0])
    for key in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline",
                "dtype", "data", "config", "roofline", "cpu_baseline", "e2e", "gpu_launches", "clocks"):
        assert key in line, key
    assert line["n_gpus"] == 1 and line["steps"] == 6 and line["warmup"] == 3 and line["value"] > 0
    assert line["config"]["workload"] and line["dtype"] == "f64" and line["gpu_launches"] > 0
    assert {"bound", "achieved", "peak", "unit", "frac", "traffic"} <= set(line["roofline"])
    assert line["e2e"]["value"] > 0 and line["e2e"]["h2d_bytes_per_step"] == 2 * 864 * 24 and line["e2e"]["d2h_bytes_per_step"] == 3 * 864 * 24
    assert {"value", "unit", "cores", "kind", "sample"} <= set(line["cpu_baseline"])
    assert line["eam"]["value"] > 0 and line["eam"]["e2e"]["value"] > 0
    assert line["chain"]["value"] > 0 and line["chain"]["neigh"]["builds"] >= 1
    # parity at the sizes the numbers are quoted on: the unmodified reference's thermo on the same input, step 0 and
    # after the parity steps, and its rebuild count (here 6^3 cells; 16.4 M / 4 M atoms on the GPU box)
    for blk in (line["parity"], line["eam"]["parity"]):
        assert blk["ok"] and blk["builds"]["equal"] and blk["steps"] == 9
        assert max(blk["pe"], blk["ke"], blk["press"]) <= 1e-6 and blk["lattice_step0"]["pe_per_atom_rel"] < 1e-9
    assert line["chain"]["parity"]["ok"] and line["parity_all_ok"]
    assert line["chain"]["cpu_baseline"]["value"] > 0


@pytest.mark.parametrize("name", ["lj2_864", "eam_500"])
def test_emu_fp64_fallback_build_keeps_the_physics(emu, name):
    """the all-FP64 list build (the fallback for boxes whose FP32 error band would be too wide) reproduces the goldens"""
    g = load_golden(name)
    md = make_device_system(emu, name, g)
    md.tune("build_fp64", 1)
    md.setup()
    a = md.gather()
    assert np.abs(a["f"] - g["a_f"]).max() <= F_RTOL * np.abs(g["a_f"]).max()
    md.run(int(g["nmore"][0]))
    np.testing.assert_allclose(thermo_vec(md.thermo_log[-1]), g["b_thermo"][:3], rtol=THERMO_RTOL)
    md.close()


def test_emu_ke_noalloc_matches_thermo(emu):
    """b200md_ke_noalloc (fixed file-scope scratch array, no allocation) against b200md_thermo's kinetic energy, two
    calls in a row with the larger system first; the size limit is refused with a message.  Under the emulator a
    device pointer is a host pointer, so numpy arrays stand in for the caller's device buffers."""
    L = emu.lib.load()
    for name in ("lj2_864", "lj_500"):
        g = load_golden(name)
        md = make_device_system(emu, name, g)
        md.setup()
        md.run(3)
        ke = md.thermo()["ke"]
        a = md.gather(sort_by_tag=False)
        v = np.ascontiguousarray(a["v"].T)
        typ, mass, out = a["type"].astype(np.int32), np.asarray(md.mass, dtype=np.float64), np.full(1, -1.0)
        p = emu.lib.ptr
        assert L.b200md_ke_noalloc(0, v.shape[1], p(v[0]), p(v[1]), p(v[2]), p(typ), p(mass), p(out), None) == 0
        np.testing.assert_allclose(0.5 * out[0], ke, rtol=1e-12)
        md.close()
    nmax = L.b200md_ke_noalloc_max_atoms()
    assert L.b200md_ke_noalloc(0, nmax + 1, p(v[0]), p(v[1]), p(v[2]), p(typ), p(mass), p(out), None) != 0
    assert "exceed" in L.b200md_last_error(None).decode()
