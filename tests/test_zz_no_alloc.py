"""b200md_ke_noalloc (include/b200md.h): the kinetic-energy sum of b200md_thermo with its per-block partials in a
fixed file-scope device array -- no allocation inside the call.  No counterpart in the reference.
Checked through the C ABI against the existing entry point (b200md_thermo) on two calls in a
row.  The file sorts last on purpose; it passes alone in a fresh process and after
the other tests (it shares no state with them: its own contexts, its own torch buffers)."""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest

from conftest import CASES, load_golden, make_device_system

KE_RTOL = 1e-12   # same atoms in the same order and block shape: the two paths differ at most by rounding order


def _ke_noalloc(pkg, torch, md, stream=None):
    """sum m v^2 of md's owned atoms through b200md_ke_noalloc, from caller-owned device buffers"""
    L = pkg.lib.load()
    a = md.gather(sort_by_tag=False)   # device order: the order b200md_thermo sums in
    dev = torch.device("cuda:0")
    v = torch.from_numpy(np.ascontiguousarray(a["v"].T)).to(dev)   # [3][n]
    typ = torch.from_numpy(a["type"].astype(np.int32)).to(dev)
    mass = torch.from_numpy(np.asarray(md.mass, dtype=np.float64)).to(dev)
    out = torch.full((1,), -1.0, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    n = v.shape[1]
    rc = L.b200md_ke_noalloc(0, n, C.c_void_p(v[0].data_ptr()), C.c_void_p(v[1].data_ptr()), C.c_void_p(v[2].data_ptr()),
                             C.c_void_p(typ.data_ptr()), C.c_void_p(mass.data_ptr()), C.c_void_p(out.data_ptr()), stream)
    assert rc == 0, L.b200md_last_error(None).decode()
    torch.cuda.synchronize()
    free1 = torch.cuda.mem_get_info()[0]
    return float(out.cpu()[0]), free0 - free1


@pytest.mark.gpu
def test_ke_noalloc_matches_thermo_twice(pkg):
    import torch

    from importlib import import_module

    units = import_module("lammps-fugaku-threadpool_b200.inputs").UNITS
    used = []
    # two calls in a row on different systems, the larger first: the second call finds the first one's partials in
    # the shared array beyond its own block count and must not read them
    for name in ("lj2_864", "lj_500"):
        g = load_golden(name)
        md = make_device_system(pkg, name, g)
        md.setup()
        md.run(5)
        ke = md.thermo()["ke"]
        s, grew = _ke_noalloc(pkg, torch, md)
        used.append(grew)
        got = 0.5 * units[CASES[name]["units"]]["mvv2e"] * s
        assert ke > 0
        np.testing.assert_allclose(got, ke, rtol=KE_RTOL)
        md.close()
    assert used[1] == 0, f"device memory in use grew by {used[1]} bytes across the second call"


@pytest.mark.gpu
def test_ke_noalloc_limits(pkg):
    import torch

    L = pkg.lib.load()
    nmax = L.b200md_ke_noalloc_max_atoms()
    assert nmax == 65536 * 256
    out = torch.zeros(1, dtype=torch.float64, device="cuda:0")
    one = torch.ones(4, dtype=torch.float64, device="cuda:0")
    typ = torch.ones(4, dtype=torch.int32, device="cuda:0")
    p = lambda t: C.c_void_p(t.data_ptr())
    # beyond the fixed array: refused before anything is launched, message names the limit
    rc = L.b200md_ke_noalloc(0, nmax + 1, p(one), p(one), p(one), p(typ), p(one), p(out), None)
    assert rc != 0
    msg = L.b200md_last_error(None).decode()
    assert "exceed" in msg and str(nmax) in msg
    assert L.b200md_ke_noalloc(0, 4, None, p(one), p(one), p(typ), p(one), p(out), None) != 0
    # an empty system sums to zero (and clears what the caller's output held)
    out.fill_(7.0)
    assert L.b200md_ke_noalloc(0, 0, None, None, None, None, None, p(out), None) == 0
    torch.cuda.synchronize()
    assert float(out.cpu()[0]) == 0.0
    # 4 atoms, mass[1] = 1, v = (1,1,1): 12
    assert L.b200md_ke_noalloc(0, 4, p(one), p(one), p(one), p(typ), p(one), p(out), None) == 0
    torch.cuda.synchronize()
    assert float(out.cpu()[0]) == 12.0
