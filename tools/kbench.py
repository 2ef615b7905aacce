#!/usr/bin/env python
"""Kernel-level timing harness (development tool, GPU box): per-kernel CUDA-event times of the pair / neighbor
kernels on a thermalised LJ or EAM system.   python tools/kbench.py [lj|eam] [cells] [label]"""
import ctypes as C
import importlib
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import bench  # noqa: E402

kind = sys.argv[1] if len(sys.argv) > 1 else "lj"
cells = int(sys.argv[2]) if len(sys.argv) > 2 else 100
label = sys.argv[3] if len(sys.argv) > 3 else ""
pkg = importlib.import_module("lammps-fugaku-threadpool_b200")
x, v, tag, box, n = bench.make_atoms(pkg, kind, cells, (1, 1, 1), (0, 0, 0), lambda a: a)
md = bench.make_system(pkg, kind, 0, 1, 0, None, (1, 1, 1))
md.box([0, 0, 0], box)
md.create_atoms(x, v, np.ones(len(x), dtype=np.int32), tag)
for kv in (sys.argv[4].split(",") if len(sys.argv) > 4 else []):   # e.g. overlap=1,build_gb=4
    k, val = kv.split("=")
    md.tune(k, int(val))
md.setup()
md.ctx.call("b200md_run", 60, md.dt, 0)  # thermalise: melt / heat the lattice
md.ctx.call("b200md_profile", 1)
bench.kernel_times(md, reset=True)
md.timers(reset=True)
t = time.perf_counter()
steps = 40
md.ctx.call("b200md_run", steps, md.dt, 0)
wall = (time.perf_counter() - t) * 1e3
ms, cnt = bench.kernel_times(md, reset=True)
names = ["pair", "eam_rho", "neigh", "nve", "halo", "sort"]
per = {k: (ms[i] / cnt[i] if cnt[i] else 0.0) for i, k in enumerate(names)}
st = md.neigh_stats()
nn = md.neighbor_list()[0]
print(f"   list: {nn.mean():.1f} neighbors per atom (full list)")
print(f"[kbench {label}] {kind} {n} atoms: step {wall / steps:.3f} ms | " + " ".join(f"{k} {per[k]:.3f}ms x{cnt[i]}" for i, k in enumerate(names) if cnt[i])
      + f" | maxneigh {st['maxneigh']} launches {md.timers()['launches']}")
for name, b in (("pair", bench.BYTES_LJ_PAIR if kind == "lj" else bench.BYTES_EAM_FORCE),):
    if per[name]:
        print(f"   {name}: {b * n / per[name] / 1e6:.0f} GB/s algorithmic = {b * n / per[name] / 1e6 / 6548.8:.3f} of measured HBM peak; "
              f"{n / per[name] / 1e3:.0f} Matom/s")
# rebuild cost: wall time of b200md_neigh_build (all kernels + host gaps) against the timed kernel classes
import torch
bench.kernel_times(md, reset=True)
torch.cuda.synchronize()
t = time.perf_counter()
for _ in range(3):
    md.ctx.call("b200md_neigh_build")
torch.cuda.synchronize()
wall = (time.perf_counter() - t) * 1e3 / 3
ms, cnt = bench.kernel_times(md, reset=True)
print(f"   rebuild: wall {wall:.3f} ms each; timed kernels: " + " ".join(f"{k} {ms[i] / 3:.3f}" for i, k in enumerate(names) if cnt[i]))
md.close()
