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
