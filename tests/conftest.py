"""Shared fixtures.  `-m "not gpu"` runs here on CPU (oracle vs golden vectors, host logic, C-ABI symbols,
gloo world_size-2 decomposition logic); `-m gpu` are the parity tests proper and call through the C ABI."""
from __future__ import annotations

import importlib
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

GOLDEN = ROOT / "tests" / "golden"
POT = ROOT / "lammps-fugaku-threadpool_b200" / "data" / "Cu_u3.eam"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def pytest_collection_modifyitems(config, items):
    """A GPU test that hangs (a kernel that never ends) must end the run, not sit there until the box is taken away:
    pytest-timeout's thread method dumps the stacks and exits the process after 10 minutes in one test."""
    if not config.pluginmanager.hasplugin("timeout"):
        return
    for item in items:
        if item.get_closest_marker("gpu") and not item.get_closest_marker("timeout"):
            item.add_marker(pytest.mark.timeout(600, method="thread"))


@pytest.fixture(scope="session")
def pkg():
    """The product package (hyphenated directory name => importlib)."""
    return importlib.import_module("lammps-fugaku-threadpool_b200")


@pytest.fixture(scope="session")
def oracle():
    import oracle as o  # test infrastructure

    o.build()
    return o


def load_golden(name: str) -> dict:
    d = np.load(GOLDEN / f"{name}.npz")
    return {k: d[k] for k in d.files}


# script settings of the golden cases (tests/golden/make_golden.py)
CASES = {
    "lj_500": dict(units="lj", ntypes=1, mass={1: 1.0}, skin=0.3, every=20, delay=0, check=0, dt=0.005, style="lj/cut",
                   cut_global=2.5, coeffs=[(1, 1, 1.0, 1.0, 2.5)], shift=False, mix="geometric"),
    "lj2_864": dict(units="lj", ntypes=2, mass={1: 1.0, 2: 1.7}, skin=0.3, every=1, delay=0, check=1, dt=0.005,
                    style="lj/cut", cut_global=2.5, coeffs=[(1, 1, 1.0, 1.0, 2.5), (2, 2, 0.8, 1.1, 2.2)], shift=True,
                    mix="arithmetic"),
    "eam_500": dict(units="metal", ntypes=1, mass={}, skin=1.0, every=1, delay=5, check=1, dt=0.005, style="eam"),
}


def thermo_vec(t: dict) -> np.ndarray:
    """[pe, ke, press] in the order the fixtures store them."""
    return np.array([t["pe"], t["ke"], t["press"]])


def lj_mixed_tables(case):
    """epsilon/sigma/cut (ntypes+1)^2 with the reference's mixing applied (src/pair.cpp:215-301 mix_energy/mix_distance)."""
    import math

    n = case["ntypes"] + 1
    eps, sig, cut = np.zeros((n, n)), np.zeros((n, n)), np.zeros((n, n))
    setf = np.zeros((n, n), dtype=bool)
    for (i, j, e, s, c) in case["coeffs"]:
        eps[i, j], sig[i, j], cut[i, j] = e, s, c
        setf[i, j] = True
    md = (lambda a, b: math.sqrt(a * b)) if case["mix"] == "geometric" else (lambda a, b: 0.5 * (a + b))
    for i in range(1, n):
        for j in range(i, n):
            if not setf[i, j]:
                eps[i, j] = math.sqrt(eps[i, i] * eps[j, j])
                sig[i, j] = md(sig[i, i], sig[j, j])
                cut[i, j] = md(cut[i, i], cut[j, j])
            eps[j, i], sig[j, i], cut[j, i] = eps[i, j], sig[i, j], cut[i, j]
    return eps, sig, cut


def make_oracle_system(oracle, name, g, state="a"):
    """OracleSystem loaded with golden state `state` of case `name`."""
    case = CASES[name]
    s = oracle.OracleSystem(case["ntypes"], case["units"])
    s.set_dt(case["dt"])
    s.set_box(g["boxlo"], g["boxhi"])
    for t, m in case["mass"].items():
        s.set_mass(t, m)
    s.set_neighbor(case["skin"], every=case["every"], delay=case["delay"], check=case["check"])
    if case["style"] == "lj/cut":
        eps, sig, cut = lj_mixed_tables(case)
        s.pair_lj_cut(eps, sig, cut, offset_flag=case["shift"])
    else:
        s.pair_eam({1: str(POT)})
    s.set_atoms(g[f"{state}_x"], g[f"{state}_v"], g[f"{state}_type"], g[f"{state}_tag"])
    return s


def make_device_system(pkg, name, g, state="a", **kw):
    """B200MD (product path over the C ABI) loaded with golden state `state` of case `name`."""
    case = CASES[name]
    md = pkg.B200MD(units=case["units"], ntypes=case["ntypes"], **kw)
    md.box(g["boxlo"], g["boxhi"])
    for t, m in case["mass"].items():
        md.set_mass(t, m)
    if case["style"] == "lj/cut":
        p = md.pair_style("lj/cut", case["cut_global"])
        for (i, j, e, s, c) in case["coeffs"]:
            p.coeff(i, j, e, s, c)
        p.modify(shift=case["shift"], mix=case["mix"])
    else:
        p = md.pair_style("eam")
        p.coeff(1, 1, str(POT))
    md.neighbor.set(case["skin"], "bin")
    md.neighbor.modify(every=case["every"], delay=case["delay"], check=case["check"])
    md.timestep(case["dt"])
    md.create_atoms(g[f"{state}_x"], g[f"{state}_v"], g[f"{state}_type"], g[f"{state}_tag"])
    return md
