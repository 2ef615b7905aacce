"""Mint the golden fixtures of tests/golden/ from the UNMODIFIED reference built by oracle/Makefile.ref
(oracle/_ref/ref_extract, a driver over the reference's C library API).  Run in the build container only
(/root/reference must exist); the resulting .npz files are committed and are what the GPU box checks against.

    python tests/golden/make_golden.py

Each case runs `run NWARM` (thermalise away from the perfect lattice, where forces are ~1e-13 noise) then
`run 0` (fresh neighbor list + forces at the current positions) -> state "a"; and the same followed by
`run NMORE` -> state "b" (positions, velocities, thermo after NMORE more steps, with reneighboring).
"""
from __future__ import annotations

import subprocess
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
import oracle  # noqa: E402

HERE = Path(__file__).resolve().parent
POT = ROOT / "lammps-fugaku-threadpool_b200" / "data" / "Cu_u3.eam"

LJ = """units lj
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
fix 1 all nve
thermo_modify norm no
"""

LJ2 = """units lj
atom_style atomic
lattice fcc 0.8442
region box block 0 {n} 0 {n} 0 {n}
create_box 2 box
create_atoms 1 box
set type 1 type/fraction 2 0.3 12345
mass 1 1.0
mass 2 1.7
velocity all create 1.2 4928459 loop geom
pair_style lj/cut 2.5
pair_coeff 1 1 1.0 1.0 2.5
pair_coeff 2 2 0.8 1.1 2.2
pair_modify shift yes mix arithmetic
neighbor 0.3 bin
neigh_modify delay 0 every 1 check yes
fix 1 all nve
thermo_modify norm no
"""

EAM = """units metal
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
fix 1 all nve
timestep 0.005
"""

CASES = {
    # name: (script, pair style, cells, nwarm, nmore)
    "lj_500": (LJ, "lj/cut", 5, 60, 40),
    "lj2_864": (LJ2, "lj/cut", 6, 60, 25),
    "eam_500": (EAM, "eam", 5, 60, 30),
}


def run_case(name, script, pstyle, n, nwarm, nmore, tmp):
    out = {}
    for state, tail in (("a", f"run {nwarm}\nrun 0\n"), ("b", f"run {nwarm}\nrun 0\nrun {nmore}\n")):
        inp = Path(tmp) / f"in.{name}.{state}"
        inp.write_text(script.format(n=n, pot=POT) + tail)
        pre = Path(tmp) / f"{name}.{state}"
        r = oracle.run_ref_extract(pre, pstyle, ["-in", str(inp), "-log", "none", "-screen", "none"], cwd=tmp)
        o = np.argsort(r["tag"])
        for k in ("tag", "type", "x", "v", "f"):
            out[f"{state}_{k}"] = r[k][o]
        m = r["meta"]
        out[f"{state}_thermo"] = np.array([m["pe"], m["ke"], m["press"], m["temp"], m["pxx"], m["pyy"], m["pzz"], m["pxy"],
                                           m["pxz"], m["pyz"]])
        out[f"{state}_nghost"] = np.array([int(m["nghost"])])
        if state == "a":
            p = r["pairs"]
            out["boxlo"] = np.array(m["boxlo"])
            out["boxhi"] = np.array(m["boxhi"])
            # undirected pairs as sorted (min_tag, max_tag) rows; image codes dropped (boxes here are > 2 cutoffs wide)
            u = np.sort(p[:, :2], axis=1)
            u = u[np.lexsort((u[:, 1], u[:, 0]))]
            out["a_pairs"] = u.astype(np.int32)
    out["nwarm"] = np.array([nwarm])
    out["nmore"] = np.array([nmore])
    np.savez_compressed(HERE / f"{name}.npz", **out)
    print(name, "atoms", len(out["a_tag"]), "pairs", len(out["a_pairs"]), "thermo a", out["a_thermo"][:3])


def main():
    if not oracle.build_ref():
        raise SystemExit("reference not available: cannot mint fixtures here")
    with tempfile.TemporaryDirectory() as tmp:
        for name, (script, pstyle, n, nwarm, nmore) in CASES.items():
            run_case(name, script, pstyle, n, nwarm, nmore, tmp)


if __name__ == "__main__":
    main()
