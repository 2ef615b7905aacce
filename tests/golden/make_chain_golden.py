#!/usr/bin/env python
"""Mint tests/golden/chain_32k.npz from the UNMODIFIED reference (oracle/_ref/ref_extract) running bench/in.chain:
the input state of bench/data.chain (positions, velocities, bonds), and the reference's results at full precision
after `run 0` (state a) and after `run 100` (state b, bench/in.chain verbatim: fix nve + fix langevin seed 904297,
neigh_modify every 1 delay 1).  Per-atom arrays of the results are kept for every 16th atom tag only (fixture
size); thermo scalars are kept in full.  Run here, where /root/reference exists:

    python tests/golden/make_chain_golden.py
"""
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
import oracle  # noqa: E402

REF = Path("/root/reference/bench")
OUT = Path(__file__).resolve().parent / "chain_32k.npz"

IN_CHAIN = """units           lj
atom_style      bond
special_bonds   fene
read_data       {data}
neighbor        0.4 bin
neigh_modify    every 1 delay 1
bond_style      fene
bond_coeff      1 30.0 1.5 1.0 1.0
pair_style      lj/cut 1.12
pair_modify     shift yes
pair_coeff      1 1 1.0 1.0 1.12
fix             1 all nve
fix             2 all langevin 1.0 1.0 10.0 904297
thermo_modify   norm no
timestep        0.012
run             {n}
"""


def read_data_chain(path):
    lines = path.read_text().splitlines()
    natoms = int([l for l in lines if l.endswith(" atoms")][0].split()[0])
    nbonds = int([l for l in lines if l.endswith(" bonds")][0].split()[0])
    lo, hi = np.zeros(3), np.zeros(3)
    for k, key in enumerate(("xlo xhi", "ylo yhi", "zlo zhi")):
        t = [l for l in lines if l.endswith(key)][0].split()
        lo[k], hi[k] = float(t[0]), float(t[1])

    def sec(name, count):
        i = [k for k, l in enumerate(lines) if l.strip() == name][0] + 2
        return [lines[i + k].split() for k in range(count)]

    at, ve, bo = sec("Atoms", natoms), sec("Velocities", natoms), sec("Bonds", nbonds)
    tag = np.array([int(t[0]) for t in at], dtype=np.int32)
    assert np.array_equal(tag, np.array([int(t[0]) for t in ve]))
    return dict(tag=tag, mol=np.array([int(t[1]) for t in at], dtype=np.int32), type=np.array([int(t[2]) for t in at], dtype=np.int32),
                x=np.array([[float(q) for q in t[3:6]] for t in at]), image=np.array([[int(q) for q in t[6:9]] for t in at], dtype=np.int32),
                v=np.array([[float(q) for q in t[1:4]] for t in ve]),
                bonds=np.array([[int(t[1]), int(t[2]), int(t[3])] for t in bo], dtype=np.int32), boxlo=lo, boxhi=hi)


def main():
    d = read_data_chain(REF / "data.chain")
    out = {f"in_{k}": v for k, v in d.items()}
    with tempfile.TemporaryDirectory() as tmp:
        for state, n in (("a", 0), ("b", 100)):
            inp = Path(tmp) / f"in.chain.{state}"
            inp.write_text(IN_CHAIN.format(data=REF / "data.chain", n=n))
            r = oracle.run_ref_extract(Path(tmp) / f"chain.{state}", "lj/cut", ["-in", str(inp), "-log", "none", "-screen", "none"], cwd=tmp)
            o = np.argsort(r["tag"])
            keep = o[::16]
            out[f"{state}_tag"] = r["tag"][keep]
            for k in ("x", "v", "f"):
                out[f"{state}_{k}"] = r[k][keep]
            m = r["meta"]
            out[f"{state}_thermo"] = np.array([m["evdwl"], m["ebond"], m["ke"], m["temp"], m["press"]])
            out[f"{state}_npairs"] = np.array([int(m["npairs"])])
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, OUT.stat().st_size, "bytes")


if __name__ == "__main__":
    main()
