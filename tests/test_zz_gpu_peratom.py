"""Per-atom energy / virial pass (b200md_pair_peratom: k_lj_peratom, k_eam_peratom) through the C ABI: its sums must be
the global tallies of the force kernels (two independent code paths over the same list), and a rigid translation of the
whole system must not change any atom's value.  The comparison with the reference's compute pe/atom and stress/atom is
tests/test_lammps_host.py::test_peratom_energy_and_stress_match_reference.  Sorts last (first hardware run)."""
from __future__ import annotations

import numpy as np
import pytest

from conftest import CASES, load_golden, make_device_system

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", list(CASES))
def test_peratom_sums_are_the_global_tallies(pkg, name):
    g = load_golden(name)
    md = make_device_system(pkg, name, g)
    md.setup()
    t = md.thermo_log[-1]
    p = md.peratom()
    assert np.array_equal(p["tag"], g["a_tag"])
    assert abs(p["eatom"].sum() - t["pe"]) <= 1e-12 * abs(t["pe"])
    terms = np.abs(p["vatom"]).sum(axis=0).max()
    assert np.abs(p["vatom"].sum(axis=0) - t["virial"]).max() <= 1e-12 * terms
    md.run(int(g["nmore"][0]))
    t = md.thermo_log[-1]
    p2 = md.peratom()
    assert abs(p2["eatom"].sum() - t["pe"]) <= 1e-12 * abs(t["pe"])
    assert not np.array_equal(p["eatom"], p2["eatom"])
    md.close()


def test_peratom_needs_a_list(pkg):
    g = load_golden("lj_500")
    md = make_device_system(pkg, "lj_500", g)
    with pytest.raises(pkg.lib.B200MDError, match="before the first neighbor build"):
        md.peratom()
    with pytest.raises(pkg.lib.B200MDError, match="pair_peratom"):
        md.ctx.call("b200md_peratom_download", None, None)
    md.close()
