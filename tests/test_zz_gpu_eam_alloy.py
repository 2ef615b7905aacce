"""pair_style eam/alloy (setfl) and eam/fs through the C ABI: the library's own file readers
(b200md_eam_init_setfl) and the multi-type EAM kernels against the oracle, whose readers are pinned to the reference's
unit-test YAML goldens (tests/test_oracle_golden.py::test_oracle_reference_unittest_eam_alloy_fs_yaml).  Sorts last: written
after the round's GPU budget was spent (run under the CPU emulation only, tests/test_emu_logic.py)."""
from __future__ import annotations

import lzma

import numpy as np
import pytest

from conftest import GOLDEN, thermo_vec

pytestmark = pytest.mark.gpu
F_RTOL = 1e-10
THERMO_RTOL = 1e-6


def alloy_inputs(pkg, n, a, seed):
    """n^3 fcc cells, two atom types scattered at random, thermal velocities: the same arrays for both sides"""
    inp = pkg.inputs
    x, cell = inp.fcc_lattice(n, n, n, a)
    tag = inp.fcc_tags(cell, n, n)
    rng = np.random.default_rng(seed)
    typ = rng.integers(1, 3, size=len(x)).astype(np.int32)
    x = x + rng.normal(scale=0.05, size=x.shape)
    v = rng.normal(scale=3.0, size=x.shape)
    v -= v.mean(axis=0)
    return x, v, typ, tag, np.array([n * a] * 3)


def potential(kind, tmp_path):
    if kind == "alloy":
        return GOLDEN / "CuNi.eam.alloy", ["Cu", "Ni"], 3.6
    p = tmp_path / "AlFe_mm.eam.fs"
    p.write_bytes(lzma.decompress((GOLDEN / "AlFe_mm.eam.fs.xz").read_bytes()))
    return p, ["Al", "Fe"], 3.9


@pytest.mark.parametrize("kind,units", [("alloy", "metal"), ("alloy", "real"), ("fs", "metal"), ("fs", "real")])
def test_eam_alloy_fs_vs_oracle(pkg, oracle, kind, units, tmp_path):
    """forces <= 1e-10 relative and thermo at step 0; PE / KE / P <= 1e-6 and the same number of rebuilds after 40 steps
    (every 1 delay 5 check yes); masses come from the file; real units exercise the readers' unit conversion"""
    file, elements, a = potential(kind, tmp_path)
    x, v, typ, tag, box = alloy_inputs(pkg, 4, a, 7)
    dt = 0.002 if units == "metal" else 2.0
    if units == "real":
        v = v * 1e-3          # A/fs instead of A/ps
    md = pkg.B200MD(units=units, ntypes=2)
    md.box([0, 0, 0], box)
    p = md.pair_style("eam/fs" if kind == "fs" else "eam/alloy")
    p.coeff("*", "*", file, *elements)
    md.neighbor.set(1.0, "bin")
    md.neighbor.modify(every=1, delay=5, check=1)
    md.timestep(dt)
    md.create_atoms(x, v, typ, tag)

    so = oracle.OracleSystem(2, units)
    so.set_dt(dt)
    so.set_box([0, 0, 0], box)
    so.set_neighbor(1.0, every=1, delay=5, check=1)
    so.pair_eam_setfl(file, elements, fs=(kind == "fs"))
    so.set_atoms(x, v, typ, tag)

    so.setup()
    md.setup()
    assert md.mass[1] == so.mass(1) and md.mass[2] == so.mass(2) and md.mass[1] != md.mass[2]
    a_, o_ = md.gather(), so.atoms()
    oo = np.argsort(o_["tag"])
    assert np.abs(a_["f"] - o_["f"][oo]).max() <= F_RTOL * np.abs(o_["f"]).max()
    np.testing.assert_allclose(thermo_vec(md.thermo_log[-1]), thermo_vec(so.thermo()), rtol=1e-9)
    so.run(40)
    md.run(40)
    np.testing.assert_allclose(thermo_vec(md.thermo_log[-1]), thermo_vec(so.thermo()), rtol=THERMO_RTOL)
    assert md.neigh_stats()["builds"] == so.ncalls
    md.close()


def test_eam_alloy_error_paths(pkg, tmp_path):
    """the reference's messages: unknown element, wrong number of element names, missing / truncated file"""
    md = pkg.B200MD(units="metal", ntypes=2)
    md.box([0, 0, 0], [20, 20, 20])
    p = md.pair_style("eam/alloy")
    with pytest.raises(ValueError, match="Incorrect args for pair coefficients"):
        p.coeff("*", "*", GOLDEN / "CuNi.eam.alloy", "Cu")
    p.coeff("*", "*", GOLDEN / "CuNi.eam.alloy", "Cu", "Zn")
    with pytest.raises(pkg.lib.B200MDError, match="No matching element in EAM potential file"):
        p.init()
    p.coeff("*", "*", tmp_path / "nope.eam.alloy", "Cu", "Ni")
    with pytest.raises(pkg.lib.B200MDError, match="Cannot open EAM potential file"):
        p.init()
    cut = tmp_path / "cut.eam.alloy"
    cut.write_text("\n".join((GOLDEN / "CuNi.eam.alloy").read_text().splitlines()[:40]) + "\n")
    p.coeff("*", "*", cut, "Cu", "Ni")
    with pytest.raises(pkg.lib.B200MDError, match="Invalid EAM potential file"):
        p.init()
    md.close()
