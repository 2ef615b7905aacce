"""Pin the oracle (oracle/oracle.c, the CPU restatement) against fixtures minted from the UNMODIFIED
reference (tests/golden/*.npz, see make_golden.py) and against the reference's shipped benchmark logs
(bench/log.6Oct16.lj.fixed.icc.1:51-52, bench/log.6Oct16.eam.fixed.icc.1:54-56).  CPU only."""
from __future__ import annotations

import numpy as np
import pytest

from conftest import CASES, GOLDEN, POT, load_golden, make_oracle_system, thermo_vec

NAMES = list(CASES)


def _undirected(pairs3):
    u = np.sort(pairs3[:, :2], axis=1)
    return u[np.lexsort((u[:, 1], u[:, 0]))]


@pytest.mark.parametrize("name", NAMES)
def test_oracle_state_a(oracle, name):
    """Fresh neighbor list + forces at the thermalised golden positions: list bit-exact as tag pairs,
    forces <= 1e-10 of max|f|, PE/KE/P <= 1e-9 relative."""
    g = load_golden(name)
    s = make_oracle_system(oracle, name, g)
    s.setup()
    assert s.nghost == int(g["a_nghost"][0])
    assert np.array_equal(_undirected(s.pairs()), g["a_pairs"])
    a = s.atoms()
    o = np.argsort(a["tag"])
    fmax = np.abs(g["a_f"]).max()
    assert np.abs(a["f"][o] - g["a_f"]).max() <= 1e-10 * fmax
    np.testing.assert_allclose(thermo_vec(s.thermo()), g["a_thermo"][:3], rtol=1e-9)


@pytest.mark.parametrize("name", NAMES)
def test_oracle_trajectory(oracle, name):
    """nmore Verlet steps with reneighboring from state a reproduce state b (x, v and thermo)."""
    g = load_golden(name)
    s = make_oracle_system(oracle, name, g)
    s.setup()
    s.run(int(g["nmore"][0]))
    a = s.atoms()
    o = np.argsort(a["tag"])
    assert np.array_equal(a["tag"][o], g["b_tag"])
    # trajectories are chaotic: a 1-ulp difference in summation order grows ~1e-16 -> 1e-9 over tens of steps
    box = g["boxhi"] - g["boxlo"]
    dx = a["x"][o] - g["b_x"]
    dx -= box * np.round(dx / box)
    assert np.abs(dx).max() < 1e-7
    np.testing.assert_allclose(thermo_vec(s.thermo()), g["b_thermo"][:3], rtol=1e-6)


def test_oracle_bench_lj_log(oracle):
    """bench/in.lj at its shipped size for 20 steps is too slow for the suite; 32k atoms, step 0 only:
    T, E_pair, TotEng, Press printed by the reference's 2016 log (per-atom, `lj` units)."""
    s = oracle.make_lj_system(20)
    s.setup()
    t = s.thermo()
    n = 32000
    assert abs(t["temp"] - 1.44) < 5e-9
    assert abs(t["pe"] / n - (-6.7733681)) < 5e-8
    assert abs((t["pe"] + t["ke"]) / n - (-4.6134356)) < 5e-8
    assert abs(t["press"] - (-5.0197073)) < 5e-8
    b = s.bin_info()
    assert (b["nbinx"], b["nstencil"]) == (23, 62)


def test_oracle_bench_eam_log(oracle):
    s = oracle.make_eam_system(str(POT), 20)
    s.setup()
    t = s.thermo()
    assert abs(t["temp"] - 1600.0) < 5e-6
    assert abs(t["pe"] - (-113280.0)) < 0.5
    assert abs(t["pe"] + t["ke"] - (-106662.09)) < 5e-3
    assert abs(t["press"] - 18703.573) < 5e-4
    assert s.bin_info()["nbinx"] == 24


def test_oracle_eam_tables_match_funcfl_header(oracle):
    """Cu_u3.eam header: nrho = nr = 500, drho = 5.01e-4, dr = 1e-2, cut = 4.95 (bench/Cu_u3.eam:2-3); the
    re-gridded tables have nr = int(rmax/dr + 0.5) = 499 knots (src/MANYBODY/pair_eam.cpp:559-560)."""
    s = oracle.OracleSystem(1, "metal")
    s.set_neighbor(1.0)
    s.pair_eam({1: str(POT)})
    t = s.eam_tables()
    assert (t["nrho"], t["nr"]) == (499, 499)
    assert abs(t["cutmax"] - 4.95) < 1e-12
    assert abs(1.0 / t["rdr"] - 0.01) < 1e-12
    assert abs(s.mass(1) - 63.55) < 1e-12


@pytest.mark.parametrize("name,units", [("unittest_atomic_pair_eam", "metal"), ("unittest_atomic_pair_eam_real", "real")])
def test_oracle_reference_unittest_eam_yaml(oracle, name, units):
    """The reference's OWN known-answer test for pair eam (unittest/force-styles/tests/atomic-pair-eam.yaml and
    atomic-pair-eam_real.yaml, minted by tests/golden/make_unittest_golden.py): 32 atoms, two funcfl elements
    (Al, Cu), a 7 A periodic box thinner than the ghost cutoff (several image layers), newton on;
    init_* = `run 0`, run_* = 4 nve steps of 1e-4 with neigh_modify delay 2 every 2 check no.  The yaml's own
    tolerance is epsilon = 6e-12 / 5e-12 (x5 for run_forces); the restatement is held to it."""
    g = load_golden(name)
    eps = float(g["epsilon"])
    s = oracle.OracleSystem(2, units)
    s.set_dt(0.0001)
    s.set_box(g["boxlo"], g["boxhi"])
    s.set_neighbor(2.0, every=2, delay=2, check=0)
    s.pair_eam({1: str(GOLDEN / "Al_jnp.eam"), 2: str(GOLDEN / "Cu_u3_units.eam")})
    s.set_atoms(g["x"], g["v"], g["type"], g["tag"])
    s.setup()
    for phase in ("init", "run"):
        if phase == "run":
            s.run(4)
        a, t = s.atoms(), s.thermo()
        o = np.argsort(a["tag"])
        f_ref = g[f"{phase}_forces"]
        scale = np.abs(f_ref).max()
        assert np.abs(a["f"][o] - f_ref).max() <= 5 * eps * scale, phase
        assert abs(t["pe"] - float(g[f"{phase}_vdwl"])) <= eps * abs(float(g[f"{phase}_vdwl"])), phase
        assert np.abs(t["virial"] - g[f"{phase}_stress"]).max() <= eps * np.abs(g[f"{phase}_stress"]).max(), phase


def unpack_fs(tmp_path):
    import lzma

    p = tmp_path / "AlFe_mm.eam.fs"
    p.write_bytes(lzma.decompress((GOLDEN / "AlFe_mm.eam.fs.xz").read_bytes()))
    return p


@pytest.mark.parametrize("name,units,fs", [("unittest_atomic_pair_eam_alloy", "metal", False),
                                           ("unittest_atomic_pair_eam_alloy_real", "real", False),
                                           ("unittest_atomic_pair_eam_fs", "metal", True),
                                           ("unittest_atomic_pair_eam_fs_real", "real", True)])
def test_oracle_reference_unittest_eam_alloy_fs_yaml(oracle, name, units, fs, tmp_path):
    """The reference's known-answer tests for pair eam/alloy (CuNi.eam.alloy, setfl) and eam/fs (AlFe_mm.eam.fs,
    Finnis-Sinclair: the density function depends on both elements) in metal and real units
    (unittest/force-styles/tests/atomic-pair-eam_{alloy,fs}{,_real}.yaml): same 32 atoms and run protocol as the
    funcfl case above, held to the yaml's own epsilon.  The real-units cases pin the reader's unit conversion,
    including the setfl reader's habit of leaving the last knot of F(rho) and z2r unconverted."""
    g = load_golden(name)
    eps = float(g["epsilon"])
    s = oracle.OracleSystem(2, units)
    s.set_dt(0.0001)
    s.set_box(g["boxlo"], g["boxhi"])
    s.set_neighbor(2.0, every=2, delay=2, check=0)
    if fs:
        s.pair_eam_setfl(unpack_fs(tmp_path), ["Al", "Fe"], fs=True)
    else:
        s.pair_eam_setfl(GOLDEN / "CuNi.eam.alloy", ["Cu", "Ni"])
    s.set_atoms(g["x"], g["v"], g["type"], g["tag"])
    s.setup()
    for phase in ("init", "run"):
        if phase == "run":
            s.run(4)
        a, t = s.atoms(), s.thermo()
        o = np.argsort(a["tag"])
        f_ref = g[f"{phase}_forces"]
        assert np.abs(a["f"][o] - f_ref).max() <= 5 * eps * np.abs(f_ref).max(), phase
        assert abs(t["pe"] - float(g[f"{phase}_vdwl"])) <= eps * abs(float(g[f"{phase}_vdwl"])), phase
        assert np.abs(t["virial"] - g[f"{phase}_stress"]).max() <= eps * np.abs(g[f"{phase}_stress"]).max(), phase


def make_chain_oracle(oracle, g):
    """bench/in.chain on the oracle: atom_style bond, special_bonds fene, lj/cut 1.12 shifted, bond fene,
    fix nve + fix langevin 1.0 1.0 10.0 904297, neighbor 0.4 / every 1 delay 1, dt 0.012"""
    s = oracle.OracleSystem(1, "lj")
    s.set_dt(0.012)
    s.set_box(g["in_boxlo"], g["in_boxhi"])
    s.set_mass(1, 1.0)
    s.set_neighbor(0.4, every=1, delay=1, check=1)
    e = np.zeros((2, 2))
    e[1, 1] = 1.0
    c = np.zeros((2, 2))
    c[1, 1] = 1.12
    s.pair_lj_cut(e, e, c, offset_flag=True)
    s.set_atoms(g["in_x"], g["in_v"], g["in_type"], g["in_tag"])
    s.set_bonds(g["in_bonds"])
    s.bond_fene({1: (30.0, 1.5, 1.0, 1.0)})
    s.fix_langevin(1.0, 1.0, 10.0, 904297)
    s.enable_sort()
    return s


def test_oracle_bench_chain_vs_reference(oracle):
    """BASELINE configs[2], bench/in.chain (32 000-bead FENE melt): the restatement of Atom::sort, NPairHalfBinNewton
    with special-bond exclusion, BondFENE, NTopoBondAll/closest_image and FixLangevin (RanMars replayed in the
    reference's local atom order) against tests/golden/chain_32k.npz, minted from the unmodified reference by
    tests/golden/make_chain_golden.py: per-atom x, v, f (every 16th tag), E_pair, E_mol, KE, T, P at step 0 and
    after the 100 steps of the benchmark, and the shipped log bench/log.6Oct16.chain.fixed.icc.1 (8 digits)."""
    g = load_golden("chain_32k")
    n = len(g["in_tag"])
    s = make_chain_oracle(oracle, g)
    s.setup()
    log = {"a": (0.97029772, 0.44484087, 20.494523, 22.394765, 4.6721833), "b": (0.9729966, 0.4361122, 20.507698, 22.40326, 4.6548819)}
    for state in ("a", "b"):
        if state == "b":
            s.run(100)
        t, b = s.thermo(), s.thermo_bond()
        got = np.array([t["pe"], b["e_bond"], t["ke"], t["temp"], t["press"]])
        ref = g[f"{state}_thermo"]
        np.testing.assert_allclose(got, ref, rtol=1e-12)
        a = s.atoms()
        where = {int(tg): i for i, tg in enumerate(a["tag"])}
        sel = np.array([where[int(tg)] for tg in g[f"{state}_tag"]])
        for k in ("x", "v", "f"):
            assert np.abs(a[k][sel] - g[f"{state}_{k}"]).max() <= 1e-10 * np.abs(g[f"{state}_{k}"]).max(), (state, k)
        assert s.L.orc_npairs(s.h) == int(g[f"{state}_npairs"][0])
        temp, epair, emol, tot, press = log[state]  # thermo normalises by N in lj units
        assert abs(t["temp"] - temp) < 5e-8 and abs(t["pe"] / n - epair) < 5e-8 and abs(b["e_bond"] / n - emol) < 5e-6
        assert abs((t["pe"] + b["e_bond"] + t["ke"]) / n - tot) < 5e-6 and abs(t["press"] - press) < 5e-7
    assert s.ncalls == 25  # "Neighbor list builds = 25" in the shipped log
