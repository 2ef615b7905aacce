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


def test_emu_chain_100_steps(emu):
    chain.test_chain_100_steps_vs_reference_and_shipped_log(emu)


def test_emu_chain_error_paths(emu):
    chain.test_chain_error_paths(emu)


# ---- paths of the atomic hot path that were written after the GPU budget was spent (opt-in on the device) ----
import test_gpu_parity as gp  # noqa: E402


@pytest.mark.parametrize("pipeline", [False, True])
@pytest.mark.parametrize("name,steps", [("lj2_864", 25), ("eam_500", 30)])
def test_emu_host_buffer_step(emu, name, steps, pipeline, monkeypatch):
    """b200md_step_host, serial form (default) and the three-stream chunk pipeline (B200MD_HOST_PIPELINE=1; the
    emulation runs streams in issue order, so this checks its indexing and chunking, not its event protocol)"""
    gp.test_host_buffer_step_equals_device_loop(emu, name, steps, pipeline, monkeypatch)


@pytest.mark.parametrize("name,steps", [("lj_500", 45), ("eam_500", 30)])
def test_emu_overlapped_halo(emu, name, steps, monkeypatch):
    """B200MD_OVERLAP=1: interior / boundary tile split of the force kernels (tile_ghost flags from the build)"""
    gp.test_overlapped_halo_equals_plain(emu, name, steps, monkeypatch)


@pytest.mark.parametrize("name", list(CASES))
def test_emu_neighbor_list_bit_exact(emu, name, monkeypatch):
    gp.test_neighbor_list_bit_exact(emu, name)
    monkeypatch.setenv("B200MD_BUILD_STAGE", "1")   # shared-memory row staging of the list stores (opt-in)
    gp.test_neighbor_list_bit_exact(emu, name)
    monkeypatch.delenv("B200MD_BUILD_STAGE")
    monkeypatch.setenv("B200MD_BUILD_FP64", "1")    # the all-FP64 build kernel
    gp.test_neighbor_list_bit_exact(emu, name)


def test_emu_piecewise_api_and_error_paths(emu):
    gp.test_piecewise_api_equals_device_loop(emu)
    gp.test_error_paths(emu)
    gp.test_run_to_run_bitwise_reproducible(emu)


@pytest.mark.parametrize("kind,n,steps", [("lj", 12, 45), ("eam", 12, 25)])
def test_emu_overlap_with_interior_tiles(emu, oracle, kind, n, steps, monkeypatch):
    """a box wide enough (12^3 cells) that many 32-atom tiles have no ghost in their lists, so the interior part of
    the split really runs under the halo: overlap == plain, bit for bit, through rebuilds"""
    out = []
    for overlap in (False, True):
        if overlap:
            monkeypatch.setenv("B200MD_OVERLAP", "1")
        else:
            monkeypatch.delenv("B200MD_OVERLAP", raising=False)
        md, _ = gp._bench_system(emu, oracle, kind, n)
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


@pytest.mark.parametrize("name,units", [("unittest_atomic_pair_eam", "metal"), ("unittest_atomic_pair_eam_real", "real")])
def test_emu_lammps_unittest_eam_yaml(emu_lmp, name, units, tmp_path):
    """host-ghost mode on a box thinner than the cutoff (several periodic image layers made by LAMMPS), two
    elements: the reference's own known-answer file through pair eam/b200"""
    lh.test_reference_unittest_eam_yaml_through_lammps(name, units, tmp_path)


def test_emu_lammps_thin_box_refused(emu_lmp, tmp_path):
    lh.test_thin_box_is_refused_by_the_resident_loop(tmp_path)


def test_emu_lammps_chain(emu_lmp, tmp_path):
    """bench/in.chain with -sf b200 end to end through the glue: VerletB200::setup on the device, bond fene/b200,
    fix langevin/b200, energies published to the reference's computes"""
    lh.test_chain_thermo_matches_reference(tmp_path)
