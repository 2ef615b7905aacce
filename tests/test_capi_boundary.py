"""The drop-in boundary without a GPU: libb200md.so builds, loads, exports every symbol include/b200md.h
declares, refuses to work without a device (no CPU fallback), and the product never touches oracle/."""
from __future__ import annotations

import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
HEADER = ROOT / "include" / "b200md.h"
PKG_DIR = ROOT / "lammps-fugaku-threadpool_b200"


def header_symbols():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(b200md_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_exported(pkg):
    L = pkg.lib.load()
    names = header_symbols()
    assert len(names) >= 30
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/b200md.h but not exported by libb200md.so"


def test_ctypes_prototypes_cover_header(pkg):
    assert sorted(pkg.lib.PROTOTYPES) == header_symbols()


def test_no_cpu_fallback_without_device(pkg):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    with pytest.raises(pkg.lib.B200MDError) as e:
        pkg.lib.Context(0)
    assert e.value.code == 4  # B200MD_ERR_CUDA
    assert "no CPU fallback" in str(e.value)


def test_null_context_is_an_error_not_a_crash(pkg):
    L = pkg.lib.load()
    assert L.b200md_set_units(None, 1.0, 1.0, 1.0, 1.0) == 5  # B200MD_ERR_ARG
    assert L.b200md_nlocal(None) == 0
    assert L.b200md_destroy(None) == 0


def test_product_never_imports_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's cpu legs may use oracle/."""
    bad = []
    for p in list(PKG_DIR.rglob("*.py")) + list(PKG_DIR.rglob("*.cu")) + list(PKG_DIR.rglob("*.cuh")) + \
            list((ROOT / "include").rglob("*.h")):
        t = p.read_text()
        if re.search(r"^\s*(import|from)\s+oracle\b", t, flags=re.M) or "liboracle" in t or "oracle/" in t.replace(
                "nothing here touches oracle/", ""):
            bad.append(str(p))
    assert not bad, bad


def test_lj_init_one_tables_match_oracle_semantics(pkg):
    """PairLJCutB200.init mirrors PairLJCut::init_one (src/pair_lj_cut.cpp:501-557): arithmetic mixing + shift."""
    class Rec:
        def __init__(self):
            self.calls = []

        def call(self, name, *a):
            self.calls.append(name)

    class Owner:
        ntypes = 2
        ctx = Rec()

    p = pkg.PairLJCutB200(Owner())
    p.settings(2.5)
    p.coeff(1, 1, 1.0, 1.0, 2.5)
    p.coeff(2, 2, 0.8, 1.1, 2.2)
    p.modify(shift=True, mix="arithmetic")
    cut = p.init()
    t = p.tables
    e12, s12, c12 = np.sqrt(0.8), 1.05, 2.35
    assert cut[1, 2] == pytest.approx(c12, rel=1e-15)
    assert t["lj1"][1, 2] == pytest.approx(48 * e12 * s12 ** 12, rel=1e-14)
    assert t["lj4"][2, 1] == pytest.approx(4 * e12 * s12 ** 6, rel=1e-14)
    assert t["offset"][1, 2] == pytest.approx(4 * e12 * ((s12 / c12) ** 12 - (s12 / c12) ** 6), rel=1e-13)
    assert Owner.ctx.calls == ["b200md_lj_init"]
    with pytest.raises(ValueError):
        p.coeff(3, 1, 1.0, 1.0)


def test_inputs_match_oracle_lattice_and_velocities(pkg, oracle):
    """inputs.py (numpy, product side) == oracle.c (C restatement) bit for bit: lattice, tags, loop-geom velocities."""
    inp = pkg.inputs
    a = inp.lattice_scale_lj(0.8442)
    assert a == oracle.lattice_scale_lj(0.8442)
    x, cell = inp.fcc_lattice(4, 3, 5, a)
    tag = inp.fcc_tags(cell, 4, 3)
    xo, to = oracle.create_fcc(4, 3, 5, a)
    assert np.array_equal(x, xo) and np.array_equal(tag, to)
    m = np.ones(len(x))
    v = inp.velocity_create_geom(x, m, 1.44, 87287, "lj")
    vo = oracle.velocity_create_geom(x, np.ones(len(x), dtype=np.int32), np.array([0.0, 1.0]), 1.44, 87287, "lj")
    np.testing.assert_allclose(v, vo, rtol=1e-13, atol=1e-15)


def test_ranmars_stream_equals_oracle(pkg, oracle):
    """fix langevin's generator is host code: the library's RanMars restatement against the oracle's (which the
    in.chain golden pins to the reference, tests/test_oracle_golden.py) -- bit-exact, several seeds, 10^5 numbers
    (more than one wrap of the 97-entry lag table and of the c/cd/cm counter)"""
    import ctypes as C

    L, O = pkg.lib.load(), oracle.lib()
    for seed in (1, 904297, 904298, 30082, 900000000):
        a, b = np.empty(100000), np.empty(100000)
        assert L.b200md_ranmars(seed, len(a), a.ctypes.data_as(C.c_void_p)) == 0
        O.orc_ranmars(seed, len(b), b.ctypes.data_as(C.c_void_p))
        assert np.array_equal(a, b)
        assert 0.0 <= a.min() and a.max() < 1.0
    assert L.b200md_ranmars(0, 1, a.ctypes.data_as(C.c_void_p)) == 5      # "Invalid seed for Marsaglia random # generator"
    assert L.b200md_ranmars(900000001, 1, a.ctypes.data_as(C.c_void_p)) == 5


def test_philox_known_answers_and_counter_uniforms(pkg):
    """the device thermostat's counter-based generator is one __host__ __device__ function: Random123's published
    known-answer vectors for Philox4x32-10, and the moments / independence of the uniforms it feeds k_langevin_counter"""
    import ctypes as C

    L = pkg.lib.load()
    kat = [([0, 0, 0, 0], [0, 0], "6627e8d5 e169c58d bc57ac4c 9b00dbd8"),
           ([0xffffffff] * 4, [0xffffffff] * 2, "408f276d 41c83b0e a20bc7c6 6d5451fd"),
           ([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0], "d16cfe09 94fdcceb 5001e420 24126ea1")]
    for ctr, key, want in kat:
        c, k, o = np.array(ctr, dtype=np.uint32), np.array(key, dtype=np.uint32), np.zeros(4, dtype=np.uint32)
        assert L.b200md_philox4x32_10(pkg.lib.ptr(c), pkg.lib.ptr(k), pkg.lib.ptr(o)) == 0
        assert " ".join(f"{x:08x}" for x in o) == want
    u = np.zeros((20000, 3))
    row = np.zeros(3)
    for t in range(len(u)):
        assert L.b200md_langevin_counter_uniforms(904297, t + 1, 17, pkg.lib.ptr(row)) == 0
        u[t] = row
    assert 0.0 < u.min() and u.max() < 1.0
    assert abs(u.mean() - 0.5) < 4 * np.sqrt(1 / 12 / u.size)                 # 4 sigma
    assert abs(u.var() - 1 / 12) < 0.002
    assert abs(np.corrcoef(u[:, 0], u[:, 1])[0, 1]) < 0.03 and abs(np.corrcoef(u[:-1, 0], u[1:, 0])[0, 1]) < 0.03
    a, b = np.zeros(3), np.zeros(3)
    L.b200md_langevin_counter_uniforms(904297, 5, 17, pkg.lib.ptr(a))
    L.b200md_langevin_counter_uniforms(904297, 5, 18, pkg.lib.ptr(b))
    assert not np.array_equal(a, b) and np.array_equal(a, u[4])               # a function of (seed, tag, step) only
    L.b200md_langevin_counter_uniforms(904298, 5, 17, pkg.lib.ptr(b))
    assert not np.array_equal(a, b)


def test_bench_reference_arm_prints_the_contract_line():
    """bench.py --impl reference (the unmodified reference's OpenMP build on the host cores, oracle/_ref/lmp_ref): one
    JSON line with the reference-arm keys, on a tiny sample"""
    import json
    import subprocess
    import sys

    if not (ROOT / "oracle" / "_ref" / "lmp_ref").exists():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    r = subprocess.run([sys.executable, "bench.py", "--impl", "reference", "--steps", "3", "--warmup", "1", "--cells", "8", "--eam-cells", "8"],
                       cwd=ROOT, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads([l for l in r.stdout.splitlines() if l.startswith("{")][-1])
    assert line["impl"] == "reference" and line["value"] > 0 and line["unit"] == "Matom-steps/s" and line["higher_is_better"] is True
    assert line["cpu_baseline"]["kind"] == "reference" and line["cpu_baseline"]["cores"] >= 1 and line["cpu_baseline"]["value"] == line["value"]
    assert line["e2e"] == {"value": line["value"], "unit": line["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
