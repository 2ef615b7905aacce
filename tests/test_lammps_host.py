"""The drop-in boundary from the reference's side: the UNMODIFIED reference (oracle/_ref/liblammps_ref.a) hosting
the b200 style classes (lammps-fugaku-threadpool_b200/lammps/) as `lmp_b200`, driven by the reference's own input
scripts (bench/in.lj, bench/in.eam as text below) with `-sf b200`.

CPU: the host builds, the plugin DSO exports lammpsplugin_init, and without a device the run dies loudly.
GPU: thermo output of `lmp_b200 -sf b200` (device-resident loop, run_style verlet/b200, and the host-ghost mode
under the stock Verlet loop) against `lmp_ref` run side by side on the same input: PE, KE(T), P within 1e-6
relative over 100 steps (north_star's gate)."""
from __future__ import annotations

import os
import re
import subprocess
from pathlib import Path

import numpy as np
import pytest



ROOT = Path(__file__).resolve().parents[1]
REFDIR = ROOT / "oracle" / "_ref"
LMP_B200 = REFDIR / "lmp_b200"
LMP_REF = REFDIR / "lmp_ref"
PLUGIN = REFDIR / "b200plugin.so"
POT = ROOT / "lammps-fugaku-threadpool_b200" / "data" / "Cu_u3.eam"

IN_LJ = """units lj
atom_style atomic
lattice fcc 0.8442
region box block 0 {n} 0 {n} 0 {n}
create_box 1 box
create_atoms 1 box
mass 1 1.0
velocity all create 1.44 87287 loop geom
pair_style lj/cut 2.5
pair_coeff 1 1 1.0 1.0 2.5
neighbor 0.3 bin
neigh_modify delay 0 every 20 check no
{runstyle}
fix 1 all nve
thermo 50
thermo_modify format float %.15g
run 100
"""

IN_EAM = """units metal
atom_style atomic
lattice fcc 3.615
region box block 0 {n} 0 {n} 0 {n}
create_box 1 box
create_atoms 1 box
pair_style eam
pair_coeff 1 1 {pot}
velocity all create 1600.0 376847 loop geom
neighbor 1.0 bin
neigh_modify every 1 delay 5 check yes
{runstyle}
fix 1 all nve
timestep 0.005
thermo 50
thermo_modify format float %.15g
run 100
"""

STOCK_LOOP = "suffix off\nrun_style verlet\nsuffix on"


def build_host():
    if not (ROOT / "oracle" / "_ref" / "liblammps_ref.a").exists():
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    if Path("/root/reference/src").exists():
        subprocess.check_call(["make", "-s", "-f", "tests/lammps_host/Makefile", "-j8"], cwd=ROOT)
    if not LMP_B200.exists():
        pytest.skip("lmp_b200 not built")


def run_lmp(exe, text, tmp_path, args=()):
    inp = tmp_path / "in.case"
    inp.write_text(text)
    env = dict(os.environ, OMP_NUM_THREADS="1")
    return subprocess.run([str(exe), "-in", str(inp), "-log", "none", "-echo", "none", *args], cwd=tmp_path, env=env,
                          capture_output=True, text=True)


def thermo_rows(stdout):
    """rows of the `Step Temp E_pair E_mol TotEng Press` table"""
    rows, on = [], False
    for line in stdout.splitlines():
        if line.split()[:2] == ["Step", "Temp"]:
            on = True
            continue
        if on:
            if line.startswith("Loop time"):
                on = False
                continue
            try:
                rows.append([float(t) for t in line.split()])
            except ValueError:
                pass
    return np.array(rows)


def thermo_rows_generic(stdout, first, second):
    """rows of a thermo table whose header starts with `first second`"""
    rows, on = [], False
    for line in stdout.splitlines():
        if line.split()[:2] == [first, second]:
            on = True
            continue
        if on:
            try:
                rows.append([float(t) for t in line.split()])
            except ValueError:
                on = False
    return rows


def test_plugin_exports_init():
    build_host()
    out = subprocess.run(["nm", "-D", "--defined-only", str(PLUGIN)], capture_output=True, text=True, check=True).stdout
    assert re.search(r"\bT lammpsplugin_init\b", out)


def test_host_fails_loudly_without_device(tmp_path):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    build_host()
    r = run_lmp(LMP_B200, IN_LJ.format(n=5, runstyle=""), tmp_path, ["-sf", "b200"])
    assert r.returncode != 0
    assert "no CPU fallback" in (r.stdout + r.stderr)


def test_host_without_suffix_is_the_reference(tmp_path):
    """without -sf b200 the host binary is the reference: same thermo, digit for digit"""
    build_host()
    a = run_lmp(LMP_B200, IN_LJ.format(n=5, runstyle=""), tmp_path)
    b = run_lmp(LMP_REF, IN_LJ.format(n=5, runstyle=""), tmp_path)
    assert a.returncode == 0 and b.returncode == 0
    np.testing.assert_array_equal(thermo_rows(a.stdout), thermo_rows(b.stdout))


@pytest.mark.gpu
@pytest.mark.parametrize("kind,n", [("lj", 20), ("eam", 20), ("lj", 7)])
@pytest.mark.parametrize("loop", ["verlet/b200", "verlet"])
def test_thermo_matches_reference(kind, n, loop, tmp_path):
    if not LMP_B200.exists() or not LMP_REF.exists():
        pytest.skip("oracle/_ref host binaries not shipped")
    text = (IN_LJ if kind == "lj" else IN_EAM)
    ref = run_lmp(LMP_REF, text.format(n=n, pot=POT, runstyle=""), tmp_path)
    assert ref.returncode == 0, ref.stdout[-2000:]
    got = run_lmp(LMP_B200, text.format(n=n, pot=POT, runstyle=STOCK_LOOP if loop == "verlet" else ""), tmp_path,
                  ["-sf", "b200"])
    assert got.returncode == 0, got.stdout[-3000:] + got.stderr[-2000:]
    # the suffix resolved to the device styles (LAMMPS falls back to the reference's class silently otherwise, and this
    # test would compare the reference with itself): every b200 style logs one line when it initialises
    assert f"b200: pair {'lj/cut' if kind == 'lj' else 'eam'}/b200 on CUDA device" in got.stdout, got.stdout[-2000:]
    assert ("b200: run_style verlet/b200" in got.stdout) == (loop == "verlet/b200"), got.stdout[-2000:]
    assert "b200:" not in ref.stdout
    R, G = thermo_rows(ref.stdout), thermo_rows(got.stdout)
    assert R.shape == G.shape and R.shape[0] == 3, (R.shape, G.shape, got.stdout[-2000:])
    np.testing.assert_array_equal(R[:, 0], G[:, 0])  # steps 0, 50, 100
    # Temp (KE), E_pair (PE), TotEng, Press: 1e-6 relative (E_mol is 0)
    for col in (1, 2, 4, 5):
        np.testing.assert_allclose(G[:, col], R[:, col], rtol=1e-6, atol=1e-9 * np.abs(R[:, col]).max())
    # step 0 goes through the same force kernels: much tighter there
    np.testing.assert_allclose(G[0, 1:], R[0, 1:], rtol=1e-10, atol=1e-12)


# ---- the device neighbor list as the reference's C library API exposes it (src/library.h:220-224) ----
REF_EXTRACT = REFDIR / "ref_extract"
B200_EXTRACT = REFDIR / "b200_extract"


def read_extract(prefix):
    """(.meta dict, undirected canonical (tag_lo, tag_hi, image code) rows, directed row count, forces by tag)"""
    meta = {}
    for line in Path(str(prefix) + ".meta").read_text().splitlines():
        k, *v = line.split()
        meta[k] = [float(t) for t in v]
    n = int(meta["nlocal"][0])
    raw = np.fromfile(str(prefix) + ".atoms.bin", dtype=np.uint8)
    tag = raw[:4 * n].view(np.int32)
    f = raw[8 * n + 48 * n:8 * n + 72 * n].view(np.float64).reshape(n, 3)
    order = np.argsort(tag)
    t = np.fromfile(str(prefix) + ".neigh.bin", dtype=np.int32).reshape(-1, 3)
    ti, tj, code = t[:, 0].astype(np.int64), t[:, 1].astype(np.int64), t[:, 2] & 63
    s = np.stack([(code >> (2 * d) & 3) - 1 for d in range(3)], axis=1)
    # the pair seen from the other atom carries the opposite image shift
    flip = (ti > tj) | ((ti == tj) & ((s[:, 0] < 0) | ((s[:, 0] == 0) & ((s[:, 1] < 0) | ((s[:, 1] == 0) & (s[:, 2] < 0))))))
    lo, hi = np.where(flip, tj, ti), np.where(flip, ti, tj)
    s = np.where(flip[:, None], -s, s)
    key = ((lo * (1 << 31) + hi) * 64 + (s[:, 0] + 1) + 4 * (s[:, 1] + 1) + 16 * (s[:, 2] + 1))
    return meta, key, len(t), f[order]


def check_device_list_through_library_api(tmp_path, kind, n, steps):
    """`b200_extract` is the oracle's list reader (oracle/ref_extract.cpp: lammps_find_pair_neighlist,
    lammps_neighlist_num_elements, lammps_neighlist_element_neighbors) linked with the b200 style classes.  With
    B200MD_HOST_LIST=1 the b200 pair style requests a full list and npair full/bin/atomonly/b200 fills it from the device
    list: every undirected (tag, tag, image) pair of the reference's half list appears exactly twice, nothing else does,
    and the forces computed from that device list equal the reference's."""
    if not B200_EXTRACT.exists() or not REF_EXTRACT.exists():
        pytest.skip("oracle/_ref host binaries not shipped")
    text = (IN_LJ if kind == "lj" else IN_EAM).replace("run 100", f"run {steps}")
    style = "lj/cut" if kind == "lj" else "eam"
    (tmp_path / "in.ref").write_text(text.format(n=n, pot=POT, runstyle=""))
    (tmp_path / "in.b200").write_text(text.format(n=n, pot=POT, runstyle=STOCK_LOOP))
    env = dict(os.environ, OMP_NUM_THREADS="1")
    ref = subprocess.run([str(REF_EXTRACT), str(tmp_path / "ref"), style, "-in", str(tmp_path / "in.ref"), "-log", "none"],
                         cwd=tmp_path, env=env, capture_output=True, text=True)
    assert ref.returncode == 0, ref.stdout[-2000:] + ref.stderr[-2000:]
    got = subprocess.run([str(B200_EXTRACT), str(tmp_path / "b200"), style + "/b200", "-sf", "b200", "-in", str(tmp_path / "in.b200"),
                          "-log", "none"], cwd=tmp_path, env=dict(env, B200MD_HOST_LIST="1"), capture_output=True, text=True)
    assert got.returncode == 0, got.stdout[-3000:] + got.stderr[-2000:]
    assert f"b200: pair {style}/b200 on CUDA device" in got.stdout, got.stdout[-2000:]
    assert "b200: npair full/bin/atomonly/b200" in got.stdout, got.stdout[-2000:]
    mr, kr, nr, fr = read_extract(tmp_path / "ref")
    mg, kg, ng, fg = read_extract(tmp_path / "b200")
    assert nr > 0 and ng == 2 * nr            # full list = both directions of the half list
    ur, cr = np.unique(kr, return_counts=True)
    ug, cg = np.unique(kg, return_counts=True)
    assert np.all(cr == 1) and np.all(cg == 2)
    assert np.array_equal(ur, ug)
    assert np.abs(fg - fr).max() <= 1e-9 * np.abs(fr).max()
    # without the switch no host list exists (the device needs none): the reader finds nothing to read
    none = subprocess.run([str(B200_EXTRACT), str(tmp_path / "none"), style + "/b200", "-sf", "b200", "-in", str(tmp_path / "in.b200"),
                           "-log", "none"], cwd=tmp_path, env=env, capture_output=True, text=True)
    assert none.returncode == 0 and not (tmp_path / "none.neigh.bin").exists()
    assert "npairs 0" in (tmp_path / "none.meta").read_text()


@pytest.mark.gpu
@pytest.mark.parametrize("kind,n,steps", [("lj", 8, 25), ("eam", 8, 12)])
def test_device_list_through_library_api(kind, n, steps, tmp_path):
    check_device_list_through_library_api(tmp_path, kind, n, steps)


# ---- the reference's own known-answer test for pair eam, run through the reference hosting the b200 styles ----
IN_UNITTEST_EAM = """atom_style atomic
atom_modify map array
neigh_modify delay 2 every 2 check no
units {units}
timestep 0.0001
newton on on
pair_style {style}
read_data data.metal
{coeffs}
{runstyle}
thermo_style custom step pe
thermo_modify format float %.17g
run 0 post no
write_dump all custom init.txt id fx fy fz modify sort id format float %.17g
fix 1 all nve
thermo 2
run 4 post no
write_dump all custom run.txt id fx fy fz modify sort id format float %.17g
"""


FUNCFL_COEFFS = (f"pair_coeff 1 1 {ROOT / 'tests' / 'golden' / 'Al_jnp.eam'}\n"
                 f"pair_coeff 2 2 {ROOT / 'tests' / 'golden' / 'Cu_u3_units.eam'}")


def write_data_metal(g, path):
    n = len(g["tag"])
    lines = ["LAMMPS data file (tests/golden/make_unittest_golden.py <- unittest/force-styles/tests/data.metal)", "",
             f"{n} atoms", "2 atom types", ""]
    for k, ax in enumerate("xyz"):
        lines.append(f"{g['boxlo'][k]:.17g} {g['boxhi'][k]:.17g} {ax}lo {ax}hi")
    lines += ["", "Masses", "", f"1 {g['mass'][1]:.17g}", f"2 {g['mass'][2]:.17g}", "", "Atoms # atomic", ""]
    for i in range(n):
        x, im = g["x"][i], g["image"][i]
        lines.append(f"{g['tag'][i]} {g['type'][i]} {x[0]:.17g} {x[1]:.17g} {x[2]:.17g} {im[0]} {im[1]} {im[2]}")
    lines += ["", "Velocities", ""]
    for i in range(n):
        v = g["v"][i]
        lines.append(f"{g['tag'][i]} {v[0]:.17g} {v[1]:.17g} {v[2]:.17g}")
    path.write_text("\n".join(lines) + "\n")


def read_forces(path):
    rows, on = [], False
    for line in path.read_text().splitlines():
        if line.startswith("ITEM: ATOMS"):
            on = True
            continue
        if on:
            rows.append([float(t) for t in line.split()])
    a = np.array(rows)
    return a[np.argsort(a[:, 0]), 1:]


@pytest.mark.gpu
@pytest.mark.parametrize("name,units", [("unittest_atomic_pair_eam", "metal"), ("unittest_atomic_pair_eam_real", "real")])
def test_reference_unittest_eam_yaml_through_lammps(name, units, tmp_path):
    """atomic-pair-eam.yaml / atomic-pair-eam_real.yaml (the reference's golden forces and energies for pair eam,
    two elements, 7 A box with several periodic image layers) through `lmp_b200 -sf b200` under the stock Verlet
    loop: pair eam/b200 in host-ghost mode, i.e. the multi-type EAM kernels and the device list build on
    LAMMPS-made ghosts.  Tolerance: forces 1e-10 of max|f| (north_star), energy 1e-10 relative."""
    if not LMP_B200.exists():
        pytest.skip("oracle/_ref host binaries not shipped")
    g = dict(np.load(ROOT / "tests" / "golden" / f"{name}.npz"))
    write_data_metal(g, tmp_path / "data.metal")
    text = IN_UNITTEST_EAM.format(units=units, style="eam", coeffs=FUNCFL_COEFFS, runstyle=STOCK_LOOP)
    r = run_lmp(LMP_B200, text, tmp_path, ["-sf", "b200"])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "b200: pair eam/b200 on CUDA device" in r.stdout, r.stdout[-2000:]
    pe = [row[1] for row in thermo_rows_generic(r.stdout, "Step", "PotEng")]
    for phase, fn, e in (("init", "init.txt", pe[0]), ("run", "run.txt", pe[-1])):
        f = read_forces(tmp_path / fn)
        ref = g[f"{phase}_forces"]
        assert np.abs(f - ref).max() <= 1e-10 * np.abs(ref).max(), phase
        assert abs(e - float(g[f"{phase}_vdwl"])) <= 1e-10 * abs(float(g[f"{phase}_vdwl"])), phase


# ---- the reference's own known-answer test for pair lj/cut on a MOLECULAR system (special_bonds weights) ----
IN_FOURMOL = """atom_style full
atom_modify map array
neigh_modify delay 2 every 2 check no
units real
timestep 0.1
newton on on
special_bonds lj/coul 0.1 0.25 0.5
pair_style lj/cut 8.0
bond_style zero
read_data data.fourmol
bond_coeff *
{coeffs}
pair_modify mix arithmetic
pair_modify shift yes
{runstyle}
thermo_style custom step pe evdwl
thermo_modify format float %.17g
run 0 post no
write_dump all custom init.txt id fx fy fz modify sort id format float %.17g
fix 1 all nve
thermo 2
run 4 post no
write_dump all custom run.txt id fx fy fz modify sort id format float %.17g
"""


def write_data_fourmol(g, path):
    """tests/golden/unittest_mol_pair_lj_cut.npz back into a data file: the atoms, velocities and BONDS of
    unittest/force-styles/tests/data.fourmol.  Its angles / dihedrals / impropers are left out: with the default
    `special_bonds ... angle no dihedral no` the 1-2 / 1-3 / 1-4 lists follow from the bond graph alone (the reference
    run on this file reproduces the YAML's forces to 3e-16, tests/golden/make_unittest_golden.py)."""
    n, nb = len(g["tag"]), len(g["bonds"])
    L = ["LAMMPS data file (tests/golden/unittest_mol_pair_lj_cut.npz <- unittest/force-styles/tests/data.fourmol)", "",
         f"{n} atoms", f"{len(g['mass']) - 1} atom types", f"{nb} bonds", f"{int(g['bonds'][:, 0].max())} bond types", ""]
    for k, ax in enumerate("xyz"):
        L.append(f"{g['boxlo'][k]:.17g} {g['boxhi'][k]:.17g} {ax}lo {ax}hi")
    L += ["", "Masses", ""] + [f"{t} {g['mass'][t]:.17g}" for t in range(1, len(g["mass"]))] + ["", "Atoms # full", ""]
    for i in range(n):
        x, im = g["x"][i], g["image"][i]
        L.append(f"{g['tag'][i]} {g['mol'][i]} {g['type'][i]} {g['q'][i]:.17g} {x[0]:.17g} {x[1]:.17g} {x[2]:.17g} {im[0]} {im[1]} {im[2]}")
    L += ["", "Velocities", ""] + [f"{g['tag'][i]} {g['v'][i][0]:.17g} {g['v'][i][1]:.17g} {g['v'][i][2]:.17g}" for i in range(n)]
    L += ["", "Bonds", ""] + [f"{k + 1} {t} {a} {b}" for k, (t, a, b) in enumerate(g["bonds"])]
    path.write_text("\n".join(L) + "\n")


def fourmol_input(g, runstyle):
    coeffs = "\n".join(f"pair_coeff {int(a)} {int(b)} {e:.17g} {s:.17g}" for a, b, e, s in g["pair_coeff"])
    return IN_FOURMOL.format(coeffs=coeffs, runstyle=runstyle)


def test_fourmol_data_file_reproduces_the_yaml_with_the_reference(tmp_path):
    """CPU: the regenerated data file (atoms + bonds only) through the UNMODIFIED reference gives the YAML's forces and
    energies -- so the GPU-side test below needs nothing from /root/reference"""
    build_host()
    g = dict(np.load(ROOT / "tests" / "golden" / "unittest_mol_pair_lj_cut.npz"))
    write_data_fourmol(g, tmp_path / "data.fourmol")
    r = run_lmp(LMP_REF, fourmol_input(g, ""), tmp_path)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    for phase in ("init", "run"):
        f = read_forces(tmp_path / f"{phase}.txt")
        assert np.abs(f - g[f"{phase}_forces"]).max() <= 1e-13 * np.abs(g[f"{phase}_forces"]).max()


@pytest.mark.gpu
def test_reference_unittest_mol_pair_lj_cut_yaml_through_lammps(tmp_path):
    """mol-pair-lj_cut.yaml (29 atoms, 5 types, arithmetic mixing, shifted, special_bonds lj/coul 0.1 0.25 0.5, a 15 A
    box under an 8 A cutoff) through `lmp_b200 -sf b200` under the stock Verlet loop: pair lj/cut/b200 in host-ghost mode
    with the host's special-bond lists applied inside the device list build (NPair::find_special) and factor_lj in the
    multi-type force kernel.  Tolerance: forces 1e-10 of max|f| (north_star), energy 1e-10 relative."""
    if not LMP_B200.exists():
        pytest.skip("oracle/_ref host binaries not shipped")
    g = dict(np.load(ROOT / "tests" / "golden" / "unittest_mol_pair_lj_cut.npz"))
    write_data_fourmol(g, tmp_path / "data.fourmol")
    r = run_lmp(LMP_B200, fourmol_input(g, STOCK_LOOP), tmp_path, ["-sf", "b200"])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "b200: pair lj/cut/b200 on CUDA device" in r.stdout, r.stdout[-2000:]
    pe = [row[1] for row in thermo_rows_generic(r.stdout, "Step", "PotEng")]
    for phase, e in (("init", pe[0]), ("run", pe[-1])):
        f = read_forces(tmp_path / f"{phase}.txt")
        ref = g[f"{phase}_forces"]
        assert np.abs(f - ref).max() <= 1e-10 * np.abs(ref).max(), phase
        assert abs(e - float(g[f"{phase}_vdwl"])) <= 1e-10 * abs(float(g[f"{phase}_vdwl"])), phase


@pytest.mark.gpu
def test_thin_box_is_refused_by_the_resident_loop(tmp_path):
    """the device makes its ghosts in one stage, which needs sub-domains at least one ghost cutoff wide: the 7 A box
    under run_style verlet/b200 must stop with the library's domain error, not run wrong"""
    if not LMP_B200.exists():
        pytest.skip("oracle/_ref host binaries not shipped")
    g = dict(np.load(ROOT / "tests" / "golden" / "unittest_atomic_pair_eam.npz"))
    write_data_metal(g, tmp_path / "data.metal")
    text = IN_UNITTEST_EAM.format(units="metal", style="eam", coeffs=FUNCFL_COEFFS, runstyle="")
    # the resident loop wants its integration fix from the first run on
    text = text.replace("run 0 post no", "fix 0 all nve\nrun 0 post no\nunfix 0", 1)
    r = run_lmp(LMP_B200, text, tmp_path, ["-sf", "b200"])
    assert r.returncode != 0
    assert "thinner than the ghost cutoff" in (r.stdout + r.stderr)


# ---- bench/in.chain (BASELINE configs[2]) through the reference hosting the b200 styles ----
IN_CHAIN = """units lj
atom_style bond
special_bonds fene
read_data data.chain
neighbor 0.4 bin
neigh_modify every 1 delay 1
bond_style {bond}
bond_coeff 1 30.0 1.5 1.0 1.0
pair_style lj/cut 1.12
pair_modify shift yes
pair_coeff 1 1 1.0 1.0 1.12
fix 1 all {nve}
fix 2 all {langevin} 1.0 1.0 10.0 904297
thermo {thermo}
thermo_modify format float %.15g
timestep 0.012
run {steps}
"""


def write_data_chain(g, path):
    """tests/golden/chain_32k.npz (the reference's read_data of bench/data.chain, dumped by
    tests/golden/make_chain_golden.py) back into a data file, 17 significant digits"""
    n, nb = len(g["in_tag"]), len(g["in_bonds"])
    lines = ["LAMMPS data file (tests/golden/chain_32k.npz <- bench/data.chain)", "", f"{n} atoms", f"{nb} bonds", "",
             "1 atom types", "1 bond types", ""]
    for k, ax in enumerate("xyz"):
        lines.append(f"{g['in_boxlo'][k]:.17g} {g['in_boxhi'][k]:.17g} {ax}lo {ax}hi")
    lines += ["", "Masses", "", "1 1.0", "", "Atoms # bond", ""]
    for i in range(n):
        x, im = g["in_x"][i], g["in_image"][i]
        lines.append(f"{g['in_tag'][i]} {g['in_mol'][i]} {g['in_type'][i]} {x[0]:.17g} {x[1]:.17g} {x[2]:.17g} {im[0]} {im[1]} {im[2]}")
    lines += ["", "Velocities", ""]
    for i in range(n):
        v = g["in_v"][i]
        lines.append(f"{g['in_tag'][i]} {v[0]:.17g} {v[1]:.17g} {v[2]:.17g}")
    lines += ["", "Bonds", ""]
    for k, (t, a, b) in enumerate(g["in_bonds"]):
        lines.append(f"{k + 1} {t} {a} {b}")
    path.write_text("\n".join(lines) + "\n")


def test_chain_data_file_round_trip_and_b200_host_styles(tmp_path):
    """(1) the data file regenerated from the golden reproduces the golden's own reference thermo (minted from
    bench/data.chain itself) digit for digit, so GPU-box runs need no /root/reference; (2) bond fene/b200,
    fix langevin/b200 and fix nve/b200 under the stock loop with a host pair style ARE the reference's classes:
    identical thermo over 20 steps"""
    build_host()
    g = dict(np.load(ROOT / "tests" / "golden" / "chain_32k.npz"))
    write_data_chain(g, tmp_path / "data.chain")
    plain = dict(bond="fene", nve="nve", langevin="langevin")
    ref0 = run_lmp(LMP_REF, IN_CHAIN.format(thermo=100, steps=0, **plain), tmp_path)
    assert ref0.returncode == 0, ref0.stdout[-2000:]
    r0 = thermo_rows(ref0.stdout)[0]   # Step Temp E_pair E_mol TotEng Press, per atom in lj units
    n = len(g["in_tag"])
    evdwl, ebond, ke, temp, press = g["a_thermo"]
    np.testing.assert_allclose(r0[1:], [temp, evdwl / n, ebond / n, (evdwl + ebond + ke) / n, press], rtol=2e-14)
    ref = run_lmp(LMP_REF, IN_CHAIN.format(thermo=10, steps=20, **plain), tmp_path)
    got = run_lmp(LMP_B200, IN_CHAIN.format(thermo=10, steps=20, bond="fene/b200", nve="nve/b200", langevin="langevin/b200"), tmp_path)
    assert ref.returncode == 0 and got.returncode == 0, got.stdout[-2000:] + got.stderr[-2000:]
    R, G = thermo_rows(ref.stdout), thermo_rows(got.stdout)
    assert R.shape == (3, 6)
    np.testing.assert_array_equal(R, G)


def test_chain_with_suffix_fails_loudly_without_device(tmp_path):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    build_host()
    g = dict(np.load(ROOT / "tests" / "golden" / "chain_32k.npz"))
    write_data_chain(g, tmp_path / "data.chain")
    r = run_lmp(LMP_B200, IN_CHAIN.format(thermo=100, steps=100, bond="fene", nve="nve", langevin="langevin"), tmp_path, ["-sf", "b200"])
    assert r.returncode != 0
    assert "no CPU fallback" in (r.stdout + r.stderr)


@pytest.mark.gpu
def test_chain_thermo_matches_reference(tmp_path):
    """bench/in.chain as shipped with `-sf b200`: pair lj/cut/b200 + bond fene/b200 + fix nve/b200 + fix
    langevin/b200 inside run_style verlet/b200 (bonded neighbor build, FENE kernel, RanMars replay) against
    lmp_ref on the same data file: Temp, E_pair, E_mol, TotEng, Press at steps 0 and 100 within 1e-6"""
    if not LMP_B200.exists() or not LMP_REF.exists():
        pytest.skip("oracle/_ref host binaries not shipped")
    g = dict(np.load(ROOT / "tests" / "golden" / "chain_32k.npz"))
    write_data_chain(g, tmp_path / "data.chain")
    text = IN_CHAIN.format(thermo=100, steps=100, bond="fene", nve="nve", langevin="langevin")
    ref = run_lmp(LMP_REF, text, tmp_path)
    got = run_lmp(LMP_B200, text, tmp_path, ["-sf", "b200"])
    assert ref.returncode == 0, ref.stdout[-2000:]
    assert got.returncode == 0, got.stdout[-3000:] + got.stderr[-2000:]
    R, G = thermo_rows(ref.stdout), thermo_rows(got.stdout)
    assert R.shape == G.shape == (2, 6), (R.shape, G.shape, got.stdout[-2000:])
    np.testing.assert_allclose(G[:, 1:], R[:, 1:], rtol=1e-6)
    np.testing.assert_allclose(G[0, 1:], R[0, 1:], rtol=1e-10)
    assert "Neighbor list builds = 25" in got.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("name,units,style", [("unittest_atomic_pair_eam_alloy", "metal", "eam/alloy"),
                                              ("unittest_atomic_pair_eam_alloy_real", "real", "eam/alloy"),
                                              ("unittest_atomic_pair_eam_fs", "metal", "eam/fs"),
                                              ("unittest_atomic_pair_eam_fs_real", "real", "eam/fs")])
def test_reference_unittest_eam_alloy_fs_yaml_through_lammps(name, units, style, tmp_path):
    """atomic-pair-eam_alloy{,_real}.yaml and atomic-pair-eam_fs{,_real}.yaml through `lmp_b200 -sf b200` under the
    stock loop: pair eam/alloy/b200 and eam/fs/b200 = the reference's own file readers + the device compute in
    host-ghost mode (eam/fs: the density table is looked up per element PAIR).  Forces 1e-10 of max|f|, energy 1e-10."""
    import lzma

    if not LMP_B200.exists():
        pytest.skip("oracle/_ref host binaries not shipped")
    g = dict(np.load(ROOT / "tests" / "golden" / f"{name}.npz"))
    write_data_metal(g, tmp_path / "data.metal")
    if style == "eam/fs":
        (tmp_path / "AlFe_mm.eam.fs").write_bytes(lzma.decompress((ROOT / "tests" / "golden" / "AlFe_mm.eam.fs.xz").read_bytes()))
        coeffs = "pair_coeff * * AlFe_mm.eam.fs Al Fe"
    else:
        coeffs = f"pair_coeff * * {ROOT / 'tests' / 'golden' / 'CuNi.eam.alloy'} Cu Ni"
    text = IN_UNITTEST_EAM.format(units=units, style=style, coeffs=coeffs, runstyle=STOCK_LOOP)
    r = run_lmp(LMP_B200, text, tmp_path, ["-sf", "b200"])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    pe = [row[1] for row in thermo_rows_generic(r.stdout, "Step", "PotEng")]
    for phase, fn, e in (("init", "init.txt", pe[0]), ("run", "run.txt", pe[-1])):
        f = read_forces(tmp_path / fn)
        ref = g[f"{phase}_forces"]
        assert np.abs(f - ref).max() <= 1e-10 * np.abs(ref).max(), phase
        assert abs(e - float(g[f"{phase}_vdwl"])) <= 1e-10 * abs(float(g[f"{phase}_vdwl"])), phase


# ---- per-atom energy / virial (compute pe/atom, compute stress/atom) ----
PERATOM_TAIL = """compute pea all pe/atom
compute sa all stress/atom NULL pair
dump 1 all custom 20 peratom.txt id c_pea c_sa[1] c_sa[2] c_sa[3] c_sa[4] c_sa[5] c_sa[6]
dump_modify 1 sort id format float %.17g
run 20
"""


def read_dump_frames(path):
    frames, rows, on = [], [], False
    for line in path.read_text().splitlines():
        if line.startswith("ITEM: TIMESTEP"):
            if rows:
                frames.append(np.array(rows))
            rows, on = [], False
        elif line.startswith("ITEM: ATOMS"):
            on = True
        elif on:
            rows.append([float(t) for t in line.split()])
    if rows:
        frames.append(np.array(rows))
    return frames


@pytest.mark.gpu
@pytest.mark.parametrize("kind", ["lj", "eam"])
@pytest.mark.parametrize("loop", ["verlet/b200", "verlet"])
def test_peratom_energy_and_stress_match_reference(kind, loop, tmp_path):
    """compute pe/atom and compute stress/atom NULL pair dumped at steps 0 and 20 from `lmp_b200 -sf b200` (device pass
    k_lj_peratom / k_eam_peratom; resident loop: scattered back by tag, stock loop: host-ghost mode) against lmp_ref on
    the same input: 1e-10 of the largest entry at step 0 (same configuration), 1e-6 after 20 steps"""
    if not LMP_B200.exists() or not LMP_REF.exists():
        pytest.skip("oracle/_ref host binaries not shipped")
    base = (IN_LJ if kind == "lj" else IN_EAM).replace("thermo 50\nthermo_modify format float %.15g\nrun 100\n", PERATOM_TAIL)
    assert "peratom.txt" in base
    ref = run_lmp(LMP_REF, base.format(n=6, pot=POT, runstyle=""), tmp_path)
    assert ref.returncode == 0, ref.stdout[-2000:]
    R = read_dump_frames(tmp_path / "peratom.txt")
    (tmp_path / "peratom.txt").unlink()
    got = run_lmp(LMP_B200, base.format(n=6, pot=POT, runstyle=STOCK_LOOP if loop == "verlet" else ""), tmp_path, ["-sf", "b200"])
    assert got.returncode == 0, got.stdout[-3000:] + got.stderr[-2000:]
    G = read_dump_frames(tmp_path / "peratom.txt")
    assert len(R) == len(G) == 2 and R[0].shape == G[0].shape == (4 * 6 ** 3, 8)
    # On the perfect fcc lattice of step 0 the EAM virial of an atom is a sum of pair terms of about 1 eV each
    # (x nktv2p = 1.6e6 in stress/atom's units) that cancel to 2e-7 eV, so round-off shows at 1e-9 of the NET value
    # -- the reference's own `compute pressure` and the sum of its stress/atom differ that much.  The step-0 tolerance
    # is therefore taken relative to the size of the terms, not of what is left of them.
    term = {"lj": 1.0, "eam": 1.6021765e6}[kind]
    for k, rtol in ((0, 1e-10), (1, 1e-6)):
        assert np.array_equal(R[k][:, 0], G[k][:, 0])
        for cols in ([1], [2, 3, 4, 5, 6, 7]):
            scale = max(np.abs(R[k][:, cols]).max(), term if cols[0] > 1 else 0.0)
            assert np.abs(G[k][:, cols] - R[k][:, cols]).max() <= rtol * scale, (k, cols)


# ---- restart / data-file round trips through the b200 pair style (the restart and data legs of the reference's
# ---- test_pair_style.cpp: write_restart / read_restart, write_data / read_data, then the same energies) ----
IN_ROUNDTRIP = """units lj
atom_style atomic
lattice fcc 0.8442
region box block 0 5 0 5 0 5
create_box 2 box
create_atoms 1 box
set group all type/fraction 2 0.3 12345
mass 1 1.0
mass 2 1.5
velocity all create 1.44 87287 loop geom
pair_style lj/cut 2.5
pair_coeff 1 1 1.0 1.0 2.5
pair_coeff 2 2 0.8 1.1 2.5
pair_modify shift yes mix arithmetic
neighbor 0.3 bin
fix 1 all nve
thermo_style custom step pe press
thermo_modify format float %.15g
run 10
write_restart rt.restart
write_data rt.data
clear
read_restart rt.restart
fix 1 all nve
thermo_style custom step pe press
thermo_modify format float %.15g
run 0
clear
units lj
atom_style atomic
pair_style lj/cut 2.5
read_data rt.data
pair_modify shift yes mix arithmetic
neighbor 0.3 bin
fix 1 all nve
thermo_style custom step pe press
thermo_modify format float %.15g
run 0
"""


@pytest.mark.gpu
def test_restart_and_data_round_trip_under_suffix_b200(tmp_path):
    """pair lj/cut/b200 inherits PairLJCut's write_restart / read_restart / write_data (src/pair_lj_cut.cpp:563-656):
    a restart written under -sf b200 restores the b200 style with its coefficients, a data file restores them through
    `Pair Coeffs`, and both reproduce the energy and pressure the first run ended with; lmp_ref beside it gives the
    same three rows"""
    if not LMP_B200.exists() or not LMP_REF.exists():
        pytest.skip("oracle/_ref host binaries not shipped")
    ref = run_lmp(LMP_REF, IN_ROUNDTRIP, tmp_path)
    assert ref.returncode == 0, ref.stdout[-2000:]
    got = run_lmp(LMP_B200, IN_ROUNDTRIP, tmp_path, ["-sf", "b200"])
    assert got.returncode == 0, got.stdout[-3000:] + got.stderr[-2000:]
    assert "lj/cut/b200" in got.stdout     # "restoring pair style lj/cut/b200 from restart"
    R = np.array(thermo_rows_generic(ref.stdout, "Step", "PotEng"))
    G = np.array(thermo_rows_generic(got.stdout, "Step", "PotEng"))
    assert R.shape == G.shape == (4, 3), (R.shape, G.shape)
    # rows: step 0, step 10, after read_restart (step 10), after read_data (step 0 of a new run)
    np.testing.assert_allclose(G[1, 1:], G[2, 1:], rtol=1e-12)
    np.testing.assert_allclose(G[1, 1:], G[3, 1:], rtol=1e-10)
    np.testing.assert_allclose(G[:, 1:], R[:, 1:], rtol=1e-6)
    np.testing.assert_allclose(G[0, 1:], R[0, 1:], rtol=1e-10)
