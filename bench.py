#!/usr/bin/env python
"""bench.py -- Matom-steps/s of the short-range MD hot path (bin -> neighbor list -> lj/cut | eam -> nve -> halo)
on N B200s of one node, with the roofline of the dominant kernel and the reference's CPU path beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--cells C] [--no-eam]

A "step" is one Verlet timestep (one pass of the hot path over all atoms of the workload, including the
neighbor rebuilds the input's `neigh_modify` schedules).  Workload at N=1: the LJ melt of bench/in.lj scaled
to 16.384 M atoms (-var x 8 -var y 8 -var z 8; BASELINE.json configs[3]); at N>1 the same 16.384 M atoms PER GPU
(weak scaling, spatial decomposition).  The EAM Cu workload (4 M atoms per GPU, configs[4]) is measured
in the same run and reported under "eam".  Metric: Finish's atom-steps/s (src/finish.cpp:139-152) / 1e6.

Only the `cpu_baseline` leg and `--impl reference` execute anything under oracle/ (the unmodified reference
built into oracle/_ref, else the oracle port); the GPU arm has no CPU fallback.
"""
from __future__ import annotations

import argparse
import ctypes as C
import importlib
import json
import os
import re
import subprocess
import sys
import tempfile
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
POT = ROOT / "lammps-fugaku-threadpool_b200" / "data" / "Cu_u3.eam"

# algorithmic bytes per owned atom (FP64, half list as the reference builds it)
NBAR_LJ, NBAR_EAM = 37.59, 37.74
BYTES_LJ_PAIR = 52 + 4 * NBAR_LJ            # 202.4 B / atom / launch of the lj/cut force kernel
BYTES_EAM_RHO = 28 + 4 * NBAR_EAM + 8 + 16  # pass A + B fused: x_i,type + ids + rho write + fp
BYTES_EAM_FORCE = 28 + 8 + 4 * NBAR_EAM + 24
BYTES_NEIGH_CHAIN = 28 + 4 * 4.87 + 4      # neighbor build of in.chain (4.87 stored half-list neighbors per bead), per build
BYTES_CHAIN_PAIR = 52 + 4 * 4.87            # its lj/cut 1.12 force kernel
BYTES_NVE = 192.0                           # fix nve, SURVEY section 8d (initial + final integrate: x rw, v rw twice, f r twice)
BYTES_LJ_STEP = 394.0                       # pair 202 + nve 192 (no rebuild)
BYTES_EAM_STEP = 606.0


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--cells", type=int, default=160, help="fcc cells per GPU per dimension for LJ (160 -> 16.384 M atoms)")
    ap.add_argument("--eam-cells", type=int, default=100, help="fcc cells per GPU per dimension for EAM (100 -> 4 M atoms)")
    ap.add_argument("--no-eam", action="store_true")
    ap.add_argument("--chain", type=int, default=8, metavar="R",
                    help="bench/in.chain replicated R x R x R (32 000 R^3 beads; BASELINE configs[2]; one GPU only; 0 = skip)")
    ap.add_argument("--no-parity", action="store_true", help="skip the parity legs (reference thermo at the quoted sizes)")
    ap.add_argument("--parity-steps", type=int, default=0, help="steps of the full-size parity leg (default min(W+K, 50))")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--halo-put", action="store_true",
                    help="N > 1: per-step halo as the fused pack-and-put kernel over peer memory instead of NCCL send/recv")
    return ap.parse_args()


def proc_grid(n):
    """ProcMap::onelevel_grid for a cubic per-rank box (src/procmap.cpp:46): min-surface factors."""
    return {1: (1, 1, 1), 2: (2, 1, 1), 4: (2, 2, 1), 8: (2, 2, 2)}[n]


# ---------------------------------------------------------------------------------------------------------
# clocks sampler (profiling recipe: nvidia-smi during the timed region)
# ---------------------------------------------------------------------------------------------------------
class Clocks:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.p, self.f = index, None, None

    def start(self):
        try:
            self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
            self.p = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=self.f, stderr=subprocess.DEVNULL)
        except OSError:
            self.p = None

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        self.p.wait()
        self.f.flush()
        rows = [r.split(",") for r in Path(self.f.name).read_text().splitlines() if r.count(",") >= 8]
        os.unlink(self.f.name)
        sm = [float(r[1]) for r in rows if r[1].strip().replace(".", "").isdigit()]
        reasons = set()
        for r in rows:
            for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7), ("sw_power_cap", 8)):
                if r[col].strip().lower() == "active":
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": float(rows[0][2]) if rows else None,
                "power_w_max": max((float(r[3]) for r in rows), default=None), "samples": len(rows), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------------------------------------------
# workloads
# ---------------------------------------------------------------------------------------------------------
WORK = {
    "lj": dict(units="lj", mass=1.0, T=1.44, seed=87287, skin=0.3, every=20, delay=0, check=0, dt=0.005),
    "eam": dict(units="metal", mass=63.55, T=1600.0, seed=376847, skin=1.0, every=1, delay=5, check=1, dt=0.005),
}


def workload_config(cells, ngpus):
    """config of the headline line -- the SAME text from both arms (the CPU arm's bounded sample is described in its
    cpu_baseline.sample, not here)"""
    per = 4 * cells ** 3
    return {"workload": f"LJ melt {per} atoms per GPU ({per * ngpus} total; bench/in.lj scaled, fcc 0.8442, T=1.44 loop geom, "
                        f"lj/cut 2.5, skin 0.3, rebuild every 20 check no, fix nve, dt 0.005); inputs larger than L2",
            "grid": list(proc_grid(ngpus))}


def make_atoms(pkg, kind, cells, grid, loc, allsum):
    """This rank's block of the global fcc lattice (cells^3 cells per rank, grid ranks): positions, tags,
    velocities exactly as `create_atoms` + `velocity all create T seed loop geom` give them (inputs.py)."""
    inp = pkg.inputs
    w = WORK[kind]
    a = inp.lattice_scale_lj(0.8442) if kind == "lj" else 3.615
    tot = [cells * g for g in grid]
    lo = [cells * l for l in loc]
    xs, tags, vs = [], [], []
    slab = max(1, min(cells, (1 << 21) // (4 * cells * cells)))  # ~2 M atoms per chunk
    for k0 in range(0, cells, slab):
        nk = min(slab, cells - k0)
        x, cell = inp.fcc_lattice(cells, cells, nk, a, lo=(lo[0], lo[1], lo[2] + k0))
        xs.append(x)
        tags.append(inp.fcc_tags(cell, tot[0], tot[1]))
        vs.append(inp.velocity_loop_geom_raw(x, w["seed"]))
    x, tag, v = np.concatenate(xs), np.concatenate(tags), np.concatenate(vs)
    natoms_total = 4 * tot[0] * tot[1] * tot[2]
    v = inp.velocity_finish(v, np.full(len(x), w["mass"]), w["T"], w["units"], natoms_total=natoms_total, allsum=allsum)
    box = [t * a for t in tot]
    return x, v, tag, box, natoms_total


def make_system(pkg, kind, device, nranks, rank, nccl_id, grid):
    w = WORK[kind]
    md = pkg.B200MD(units=w["units"], ntypes=1, device=device, nranks=nranks, rank=rank, nccl_id=nccl_id, procgrid=grid)
    if kind == "lj":
        md.set_mass(1, 1.0)
        md.pair_style("lj/cut", 2.5).coeff(1, 1, 1.0, 1.0, 2.5)
    else:
        md.pair_style("eam").coeff(1, 1, str(POT))
    md.neighbor.set(w["skin"], "bin")
    md.neighbor.modify(every=w["every"], delay=w["delay"], check=w["check"])
    md.timestep(w["dt"])
    return md


def kernel_times(md, reset=False):
    ms = np.zeros(6)
    cnt = np.zeros(6, dtype=np.int64)
    md.ctx.call("b200md_kernel_times", ms.ctypes.data_as(C.c_void_p), cnt.ctypes.data_as(C.c_void_p), int(reset))
    return ms, cnt


# ---------------------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------------------
def time_resident(torch, dist, md, steps, warmup, device, nranks):
    """W untimed + K timed Verlet steps on device-resident atoms (b200md_run), CUDA events on the library's own
    stream, barrier + synchronize on both sides, max over ranks.  Returns (ms for the K steps, kernel launches in them)."""
    ext = torch.cuda.ExternalStream(md.ctx.L.b200md_stream(md.ctx.h), device=device)
    if warmup > 0:
        md.ctx.call("b200md_run", warmup, md.dt, 0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(device)
    if nranks > 1:
        dist.barrier()
    launches0 = md.timers()["launches"]
    with torch.cuda.stream(ext):
        e0.record()
    md.ctx.call("b200md_run", steps, md.dt, 0)
    with torch.cuda.stream(ext):
        e1.record()
    torch.cuda.synchronize(device)
    launches = md.timers()["launches"] - launches0   # this library's kernel launches INSIDE the timed region
    if nranks > 1:
        dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    if nranks > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item()), int(launches)


def numa_local_affinity(device):
    """Pin this process to the CPUs next to its GPU (nvmlDeviceGetCpuAffinity) so that the pinned host buffers of the e2e
    leg are allocated on that NUMA node and the copies do not cross the socket link -- several ranks on one host otherwise
    share whatever node the scheduler put them on.  Returns (old affinity, new affinity) or None if it cannot be done."""
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(int(device))
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        near = {64 * w + b for w, m in enumerate(words) for b in range(64) if (int(m) >> b) & 1}
        old = os.sched_getaffinity(0)
        new = near & old
        if not new or new == old:
            return None
        os.sched_setaffinity(0, new)
        return old, new
    except Exception:  # noqa: BLE001  (no NVML, no permission, cgroup without those CPUs ...)
        return None


def time_host_buffers(torch, dist, md, steps, warmup, device, nranks):
    """The same steps through the host-buffer entry point (b200md_step_host): every step copies x, v from pinned
    host arrays to the device and x, v, f back -- what a Pair/Fix plugin under the stock host-resident Verlet
    loop pays (the reference GPU package's model).  Returns (ms for K steps, h2d bytes/step, d2h bytes/step)."""
    n = md.nlocal
    cap = n + n // 64 + 1024  # room for atoms that migrate in at a rebuild (several ranks)
    aff = numa_local_affinity(device)
    time_host_buffers.numa = None if aff is None else {"cpus_before": len(aff[0]), "cpus_near_gpu": len(aff[1])}
    hx = torch.empty((cap, 3), dtype=torch.float64).pin_memory()
    hv = torch.empty((cap, 3), dtype=torch.float64).pin_memory()
    hf = torch.empty((cap, 3), dtype=torch.float64).pin_memory()
    md.ctx.call("b200md_download_atoms", C.c_void_p(hx.data_ptr()), C.c_void_p(hv.data_ptr()), None, None, None, None)
    e, vir, nout = C.c_double(0), np.zeros(6), C.c_int(0)
    args = (md.dt, 0, 0, cap, C.c_void_p(hx.data_ptr()), C.c_void_p(hv.data_ptr()), C.c_void_p(hf.data_ptr()), None,
            C.byref(nout), C.byref(e), vir.ctypes.data_as(C.c_void_p))
    for _ in range(warmup):
        md.ctx.call("b200md_step_host", *args)
    ext = torch.cuda.ExternalStream(md.ctx.L.b200md_stream(md.ctx.h), device=device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(device)
    if nranks > 1:
        dist.barrier()
    with torch.cuda.stream(ext):
        e0.record()
    t0 = time.perf_counter()
    for _ in range(steps):
        md.ctx.call("b200md_step_host", *args)
    with torch.cuda.stream(ext):
        e1.record()
    torch.cuda.synchronize(device)
    wall_ms = (time.perf_counter() - t0) * 1e3
    ms = torch.tensor([max(e0.elapsed_time(e1), wall_ms), float(2 * n * 24), float(3 * n * 24)], dtype=torch.float64, device=device)
    if nranks > 1:
        mx = ms.clone()
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        dist.all_reduce(ms, op=dist.ReduceOp.SUM)
        ms[0] = mx[0]
    if aff is not None:
        os.sched_setaffinity(0, aff[0])   # the CPU legs that follow use every core again
    return float(ms[0].item()), int(ms[1].item()), int(ms[2].item())


def run_workload(torch, dist, pkg, kind, cells, args, device, nranks, rank, nccl_id, grid, loc, allsum):
    t_gen = time.perf_counter()
    x, v, tag, box, natoms_total = make_atoms(pkg, kind, cells, grid, loc, allsum)
    t_gen = time.perf_counter() - t_gen
    md = make_system(pkg, kind, device, nranks, rank, nccl_id(), grid)
    md.box([0, 0, 0], box)
    md.create_atoms(x, v, np.ones(len(x), dtype=np.int32), tag, natoms_total=natoms_total)
    del x, v, tag
    if nranks > 1 and getattr(args, "halo_put", False):
        def allgather(blob):
            got = [None] * nranks
            dist.all_gather_object(got, blob)
            return got
        md.halo_put(max(1 << 20, md.nlocal // 4), allgather)   # landing buffers for up to a quarter of the owned atoms in ghosts
    md.setup()
    out = dict(natoms=natoms_total, atoms_per_gpu=md.nlocal, thermo0=md.thermo_log[-1])
    out["halo"] = "pack-and-put over peer memory" if (nranks > 1 and getattr(args, "halo_put", False)) else ("nccl send/recv" if nranks > 1 else "self")

    # ---- parity leg (untimed, before the warm-up): thermo at step 0 and after P steps against the unmodified reference
    # on the SAME input at the SAME size (N = 1), or against closed forms of the perfect lattice (any N)
    if not args.no_parity:
        psteps = args.parity_steps or min(args.steps + args.warmup, 50)
        md.run(psteps)
        out["parity_run"] = dict(steps=psteps, thermo=md.thermo_log[-1], builds=md.neigh_stats()["builds"])

    # ---- timed region: device-resident steps ---------------------------------------------------------
    md.timers(reset=True)
    clocks = Clocks(device)
    if rank == 0:
        clocks.start()
    ms, launches = time_resident(torch, dist, md, args.steps, args.warmup, device, nranks)
    out["clocks"] = clocks.stop() if rank == 0 else None
    out["ms"] = ms
    out["launches"] = launches
    out["value"] = natoms_total * args.steps / (ms * 1e-3) / 1e6

    # ---- second pass with per-kernel CUDA events (no host syncs added) for the roofline ---------------
    md.ctx.call("b200md_profile", 1)
    kernel_times(md, reset=True)
    ms2, _ = time_resident(torch, dist, md, args.steps, 0, device, nranks)
    kms, kcnt = kernel_times(md, reset=True)
    md.ctx.call("b200md_profile", 0)
    out["ms_profiled_pass"] = ms2
    out["kernel_ms"] = {k: (float(kms[i]), int(kcnt[i])) for i, k in enumerate(["pair", "eam_rho", "neigh", "nve", "halo", "sort"])}
    out["thermo1"] = md.thermo()
    out["neigh"] = md.neigh_stats()
    out["nghost"] = md.nghost
    out["mem_gb"] = float(md.ctx.L.b200md_memory_usage(md.ctx.h)) / 1e9

    out["gen_s"] = t_gen
    return out, md


def run_chain(torch, dist, pkg, r, args, device):
    """BASELINE configs[2]: bench/in.chain (FENE bead-spring melt, special_bonds fene, fix nve + fix langevin) replicated
    r^3 times like `replicate r r r`, on one GPU.  The thermostat draws from the counter-based device generator (the
    RanMars replay that reproduces the reference's trajectory is a one-rank parity mode with a host draw per step)."""
    g = dict(np.load(ROOT / "tests" / "golden" / "chain_32k.npz"))
    R = pkg.inputs.replicate_chain(g, r)
    md = pkg.B200MD(units="lj", ntypes=1, device=device)
    md.box(R["boxlo"], R["boxhi"])
    md.set_mass(1, 1.0)
    p = md.pair_style("lj/cut", 1.12)
    p.coeff(1, 1, 1.0, 1.0, 1.12)
    p.modify(shift=True)
    md.neighbor.set(0.4, "bin")
    md.neighbor.modify(every=1, delay=1, check=True)
    md.timestep(0.012)
    md.create_atoms(R["x"], R["v"], R["type"], R["tag"])
    md.special_bonds("fene")
    md.bonds(R["tag"], *pkg.inputs.bond_arrays(R["tag"], R["bonds"]))
    md.bond_style("fene", 1).coeff(1, 30.0, 1.5, 1.0, 1.0)
    md.fix("langevin", 1.0, 1.0, 10.0, 904297, rng="counter")
    natoms = len(R["tag"])
    del R
    md.setup()
    out = dict(natoms=natoms, atoms_per_gpu=natoms, thermo0=md.thermo_log[-1], builds0=md.neigh_stats()["builds"])
    md.timers(reset=True)
    ms, launches = time_resident(torch, dist, md, args.steps, args.warmup, device, 1)
    out["ms"] = ms
    out["value"] = natoms * args.steps / (ms * 1e-3) / 1e6
    out["launches"] = launches
    md.ctx.call("b200md_profile", 1)
    kernel_times(md, reset=True)
    time_resident(torch, dist, md, args.steps, 0, device, 1)
    kms, kcnt = kernel_times(md, reset=True)
    md.ctx.call("b200md_profile", 0)
    out["kernel_ms"] = {k: (float(kms[i]), int(kcnt[i])) for i, k in enumerate(["pair+bond", "eam_rho", "neigh", "nve", "halo", "sort"])}
    out["thermo1"] = md.thermo()
    out["neigh"] = md.neigh_stats()
    md.close()
    return out


def run_decomposed_parity(torch, dist, pkg, kind, device, nranks, rank, nccl_id, grid, loc, allsum, steps=60, cells=None):
    """N > 1: a small decomposed system (24^3 fcc cells per rank for lj, 20^3 for eam) run for `steps` steps before the
    timed workload -- migration, borders, halo and the rebuild votes all exercised -- and compared on rank 0 with the
    unmodified reference run on the host on the same global box: thermo at step 0 and `steps`, the rebuild count, and
    the atoms themselves (every tag 1..N owned exactly once: count, sum and a weighted sum of the owned tags)."""
    cells = cells or (24 if kind == "lj" else 20)
    x, v, tag, box, natoms = make_atoms(pkg, kind, cells, grid, loc, allsum)
    md = make_system(pkg, kind, device, nranks, rank, nccl_id(), grid)
    md.box([0, 0, 0], box)
    md.create_atoms(x, v, np.ones(len(x), dtype=np.int32), tag, natoms_total=natoms)
    md.setup()
    t0 = md.thermo_log[-1]
    md.run(steps)
    t1 = md.thermo_log[-1]
    builds = md.neigh_stats()["builds"]
    tg = md.gather(sort_by_tag=False)["tag"].astype(np.int64)
    chk = allsum(np.array([len(tg), tg.sum(), (tg * (tg % 4099)).sum()], dtype=np.float64))   # all exact below 2^53
    nmig = md.nlocal != len(x)
    md.close()
    allt = np.arange(1, natoms + 1, dtype=np.int64)
    tags_ok = bool(chk[0] == natoms and chk[1] == allt.sum() and chk[2] == (allt * (allt % 4099)).sum())
    if rank != 0:
        return None
    dims = tuple(cells * g for g in grid)
    ref = ref_thermo(kind, dims, steps, os.cpu_count() or 1)
    return parity_block(f"oracle/_ref/lmp_ref (unmodified reference, OPENMP styles) on the host, same {dims[0]}x{dims[1]}x{dims[2]}-cell box, "
                        f"{nranks} ranks on the device side", t0, t1, builds, ref,
                        extra={"tags": {"owned_exactly_once": tags_ok, "count": int(chk[0])}, "rank0_atom_count_changed": bool(nmig)})


def run_e2e(torch, dist, md, out, args, device, nranks):
    """End to end through host buffers: every rank moves its own atoms over its own PCIe link every step."""
    k = max(3, min(args.steps, 20))
    ems, bi, bo = time_host_buffers(torch, dist, md, k, 3, device, nranks)
    out["e2e"] = dict(value=out["natoms"] * k / (ems * 1e-3) / 1e6, unit="Matom-steps/s", h2d_bytes_per_step=bi,
                      d2h_bytes_per_step=bo, steps=k, api="b200md_step_host (pinned host x,v in; x,v,f out every step)",
                      form="three-stream pipeline", numa=getattr(time_host_buffers, "numa", None))


class Watchdog:
    """Bounds a phase that contains collectives: if one rank fails inside it the others would wait for ever, and the
    line with the device-resident numbers (already measured) would be lost.  On expiry rank 0 prints the line it has
    and every rank leaves with status 0."""

    def __init__(self, seconds, emit):
        import threading

        self.t = threading.Timer(seconds, self.fire)
        self.t.daemon = True
        self.emit = emit

    def fire(self):
        try:
            self.emit("e2e phase did not finish in time")
        finally:
            os._exit(0)

    def __enter__(self):
        self.t.start()
        return self

    def __exit__(self, *exc):
        self.t.cancel()
        return False


# ---------------------------------------------------------------------------------------------------------
# CPU arm: the unmodified reference (oracle/_ref/lmp_ref, OPENMP package) or, failing that, the oracle port
# ---------------------------------------------------------------------------------------------------------
REF_HEAD = {
    "lj": """units lj
atom_style atomic
lattice fcc 0.8442
region box block 0 {nx} 0 {ny} 0 {nz}
create_box 1 box
create_atoms 1 box
mass 1 1.0
velocity all create 1.44 87287 loop geom
pair_style lj/cut 2.5
pair_coeff 1 1 1.0 1.0 2.5
neighbor 0.3 bin
neigh_modify delay 0 every 20 check no
fix 1 all nve
""",
    "eam": """units metal
atom_style atomic
lattice fcc 3.615
region box block 0 {nx} 0 {ny} 0 {nz}
create_box 1 box
create_atoms 1 box
pair_style eam
pair_coeff 1 1 {pot}
velocity all create 1600.0 376847 loop geom
neighbor 1.0 bin
neigh_modify every 1 delay 5 check yes
fix 1 all nve
timestep 0.005
""",
    # bench/in.chain on the data file regenerated from tests/golden/chain_32k.npz, replicated like in.chain.scaled
    "chain": """units lj
atom_style bond
special_bonds fene
read_data data.chain
replicate {nx} {ny} {nz}
neighbor 0.4 bin
neigh_modify every 1 delay 1
bond_style fene
bond_coeff 1 30.0 1.5 1.0 1.0
pair_style lj/cut 1.12
pair_modify shift yes
pair_coeff 1 1 1.0 1.0 1.12
fix 1 all nve
fix 2 all langevin 1.0 1.0 10.0 904297
timestep 0.012
""",
}
REF_TIMED = "thermo 0\nrun {w}\nrun {k}\n"
# totals, 15 digits: what the parity legs compare (src/thermo.cpp:182-189 keywords; norm no = extensive values)
REF_THERMO = "thermo_style custom step temp pe ke press evdwl ebond\nthermo_modify format float %.15g norm no\nthermo {t}\nrun {p}\n"


def write_data_chain(g, path):
    """tests/golden/chain_32k.npz (the reference's own read_data of bench/data.chain, tests/golden/make_chain_golden.py)
    back into a data file, 17 significant digits -- /root/reference does not exist on the GPU box"""
    n, nb = len(g["in_tag"]), len(g["in_bonds"])
    lines = ["LAMMPS data file (tests/golden/chain_32k.npz <- bench/data.chain)", "", f"{n} atoms", f"{nb} bonds", "",
             "1 atom types", "1 bond types", ""]
    for k, ax in enumerate("xyz"):
        lines.append(f"{g['in_boxlo'][k]:.17g} {g['in_boxhi'][k]:.17g} {ax}lo {ax}hi")
    lines += ["", "Masses", "", "1 1.0", "", "Atoms # bond", ""]
    X, IM, T, M, TY = g["in_x"], g["in_image"], g["in_tag"], g["in_mol"], g["in_type"]
    lines += [f"{T[i]} {M[i]} {TY[i]} {X[i][0]:.17g} {X[i][1]:.17g} {X[i][2]:.17g} {IM[i][0]} {IM[i][1]} {IM[i][2]}" for i in range(n)]
    lines += ["", "Velocities", ""]
    V = g["in_v"]
    lines += [f"{T[i]} {V[i][0]:.17g} {V[i][1]:.17g} {V[i][2]:.17g}" for i in range(n)]
    lines += ["", "Bonds", ""]
    lines += [f"{k + 1} {t} {a} {b}" for k, (t, a, b) in enumerate(g["in_bonds"])]
    Path(path).write_text("\n".join(lines) + "\n")


def ref_launch(kind, dims, tail, threads):
    """one launch of the UNMODIFIED reference (oracle/_ref/lmp_ref, OPENMP styles when threads > 1) on the workload
    `kind` with dims = (nx, ny, nz) fcc cells (lj, eam) or replication factors (chain); returns its stdout"""
    import oracle

    with tempfile.TemporaryDirectory() as tmp:
        if kind == "chain":
            write_data_chain(dict(np.load(ROOT / "tests" / "golden" / "chain_32k.npz")), Path(tmp) / "data.chain")
        inp = Path(tmp) / "in.bench"
        inp.write_text(REF_HEAD[kind].format(nx=dims[0], ny=dims[1], nz=dims[2], pot=POT) + tail)
        cmd = [str(oracle.LMP_REF), "-in", str(inp), "-log", "none", "-echo", "none"]
        if threads > 1:
            cmd += ["-sf", "omp", "-pk", "omp", str(threads)]
        env = dict(os.environ, OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="true")
        return subprocess.run(cmd, cwd=tmp, env=env, capture_output=True, text=True, check=True).stdout


def ref_run(kind, n, steps, warmup, threads):
    """timing launch on an n^3 sample: `run W` then `run K`; returns (Matom-steps/s of the K-step loop as Finish prints
    it, loop seconds)"""
    txt = ref_launch(kind, (n, n, n), REF_TIMED.format(w=warmup, k=steps), threads)
    loops = re.findall(r"Loop time of ([0-9.eE+-]+) on \d+ procs for (\d+) steps with (\d+) atoms", txt)
    sec, st, na = loops[-1]
    return int(na) * int(st) / float(sec) / 1e6, float(sec)


def ref_thermo(kind, dims, steps, threads):
    """parity launch: the reference's own thermo (totals, 15 digits) at step 0 and at step `steps`, its neighbor-build
    count and atom count, on exactly the GPU arm's input"""
    txt = ref_launch(kind, dims, REF_THERMO.format(t=max(1, steps), p=steps), threads)
    rows, on = [], False
    for line in txt.splitlines():
        t = line.split()
        if t[:2] == ["Step", "Temp"]:
            on = True
            continue
        if on:
            try:
                rows.append([float(v) for v in t])
            except ValueError:
                on = False
    keys = ("step", "temp", "pe", "ke", "press", "evdwl", "ebond")
    m = re.search(r"Loop time of ([0-9.eE+-]+) on \d+ procs for (\d+) steps with (\d+) atoms", txt)
    b = re.search(r"Neighbor list builds = (\d+)", txt)
    d = re.search(r"Dangerous builds = (\d+)", txt)
    return dict(step0=dict(zip(keys, rows[0])), last=dict(zip(keys, rows[-1])), builds=int(b.group(1)) if b else None,
                dangerous=int(d.group(1)) if d else None, natoms=int(m.group(3)) if m else None,
                loop_s=float(m.group(1)) if m else None)


def lammps_host_leg(kind, cells, steps, warmup):
    """What a user of the drop-in sees: the UNMODIFIED reference hosting the b200 style classes (oracle/_ref/lmp_b200,
    lammps-fugaku-threadpool_b200/lammps/) on the same input script with `-sf b200` -- LAMMPS creates the atoms on the host,
    run_style verlet/b200 keeps them on the device, Finish prints the loop time of the K-step run."""
    exe = ROOT / "oracle" / "_ref" / "lmp_b200"
    if not exe.exists():
        return None
    with tempfile.TemporaryDirectory() as tmp:
        inp = Path(tmp) / "in.bench"
        inp.write_text(REF_HEAD[kind].format(nx=cells, ny=cells, nz=cells, pot=POT) + REF_TIMED.format(w=warmup, k=steps))
        t = time.perf_counter()
        r = subprocess.run([str(exe), "-in", str(inp), "-log", "none", "-echo", "none", "-sf", "b200"], cwd=tmp,
                           env=dict(os.environ, OMP_NUM_THREADS="1"), capture_output=True, text=True)
        wall = time.perf_counter() - t
    if r.returncode != 0 or "b200: run_style verlet/b200" not in r.stdout:
        return {"error": (r.stdout[-300:] + r.stderr[-300:]).strip()}
    loops = re.findall(r"Loop time of ([0-9.eE+-]+) on \d+ procs for (\d+) steps with (\d+) atoms", r.stdout)
    sec, st, na = loops[-1]
    return {"value": int(na) * int(st) / float(sec) / 1e6, "unit": "Matom-steps/s", "loop_s": float(sec), "steps": int(st), "natoms": int(na),
            "process_wall_s": wall, "api": "lmp_b200 -sf b200 -in <bench/in.%s scaled> (run_style verlet/b200, Loop time as Finish prints it)" % kind}


PARITY_TOL = 1e-6   # north_star: thermo (PE, KE, pressure) within 1e-6 relative of the reference's CPU path


def parity_block(label, got0, got1, builds, ref, extra=None):
    """relative errors of the device thermo against the reference's at step 0 and at the last parity step"""
    def rel(g, r):
        return abs(g - r) / max(abs(r), 1e-300)

    out = {"against": label, "steps": int(ref["last"]["step"]), "natoms": ref["natoms"], "tolerance": PARITY_TOL}
    worst = 0.0
    for name, g in (("step0", got0), ("last", got1)):
        r = ref[name]
        e = {k: rel(g[k], r[k]) for k in ("pe", "ke", "press")}
        if "ebond" in g:
            e["ebond"] = rel(g["ebond"], r["ebond"])
        out[name] = e
        worst = max(worst, *e.values())
    out["pe"], out["ke"], out["press"] = (max(out["step0"][k], out["last"][k]) for k in ("pe", "ke", "press"))
    out["builds"] = {"b200": int(builds), "reference": ref["builds"], "equal": int(builds) == ref["builds"]}
    out["reference_thermo_last"] = {k: ref["last"][k] for k in ("pe", "ke", "press")}
    out["b200_thermo_last"] = {k: float(got1[k]) for k in ("pe", "ke", "press")}
    out["ok"] = bool(worst <= PARITY_TOL and out["builds"]["equal"])
    if extra:
        out.update(extra)
    return out


def port_run(kind, n, steps, warmup):
    """Fallback when oracle/_ref is absent: the scalar C port (1 core)."""
    import oracle

    s = oracle.make_lj_system(n) if kind == "lj" else oracle.make_eam_system(str(POT), n)
    s.setup()
    s.run(warmup)
    t = time.perf_counter()
    s.run(steps)
    sec = time.perf_counter() - t
    return 4 * n ** 3 * steps / sec / 1e6, sec


def cpu_measure(kind, steps, warmup, budget_s, cells_cap):
    """Size an n^3 sample of the workload (fcc cells for lj / eam, replicas of the 32 000-bead melt for chain) so that
    W+K steps take about `budget_s` on this host, run it."""
    import oracle

    cores = os.cpu_count() or 1
    have_ref = oracle.ref_available()
    per_cell = 32000 if kind == "chain" else 4
    if have_ref:
        # calibration on the shipped 32k case, 10 steps: all host threads (OPENMP styles) against one thread -- the
        # reference arm gets whichever is faster here (in.chain's many short parallel regions can lose to one thread)
        n0 = 1 if kind == "chain" else 20
        rate, _ = ref_run(kind, n0, 10, 2, cores)
        if cores > 1:
            rate1, _ = ref_run(kind, n0, 10, 2, 1)
            if rate1 > rate:
                rate, cores = rate1, 1
        kindname = "reference"
    elif kind == "chain":
        return None   # the scalar port of in.chain is a parity tool, not a timing baseline
    else:
        oracle.build()
        rate, _ = port_run(kind, 10, 5, 1)
        cores, kindname = 1, "port"
    atoms = rate * 1e6 * budget_s / max(1, steps + warmup)
    n = int(max(1 if kind == "chain" else 10, min(cells_cap, round((atoms / per_cell) ** (1 / 3)))))
    val, sec = ref_run(kind, n, steps, warmup, cores) if have_ref else port_run(kind, n, steps, warmup)
    what = f"bench/in.chain replicated {n}^3 = {per_cell * n ** 3} beads" if kind == "chain" else f"bench/in.{kind} scaled to {n}^3 fcc cells = {4 * n ** 3} atoms"
    sample = (f"{what}, {steps} steps after {warmup} warm-up, "
              + ((f"oracle/_ref/lmp_ref -sf omp -pk omp {cores}" if cores > 1 else "oracle/_ref/lmp_ref, 1 thread (faster than the OPENMP styles "
                  "on this host for this input)") if have_ref else "oracle/oracle.c port, 1 thread")
              + f", loop time {sec:.2f} s")
    return dict(value=val, unit="Matom-steps/s", cores=cores, kind=kindname, sample=sample)


# ---------------------------------------------------------------------------------------------------------
def main_reference(args, rank):
    if rank != 0:
        return
    kind = "lj"
    cb = cpu_measure(kind, args.steps, args.warmup, budget_s=90.0, cells_cap=args.cells)
    line = {"impl": "reference", "metric": "Matom-steps/s (LJ melt, bench/in.lj scaled)", "value": cb["value"], "unit": "Matom-steps/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args.cells, args.gpus),
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "Matom-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    if not args.no_eam:
        line["eam"] = cpu_measure("eam", args.steps, args.warmup, budget_s=60.0, cells_cap=args.eam_cells)
    print(json.dumps(line), flush=True)


def main():
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return main_reference(args, rank)
    import torch
    import torch.distributed as dist

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the B200 arm has no CPU fallback (use --impl reference for the CPU arm)")
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("launch with: python -m torch.distributed.run --nproc-per-node N bench.py --gpus N ...")
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    pkg = importlib.import_module("lammps-fugaku-threadpool_b200")
    torch.cuda.set_device(local)
    grid = proc_grid(world)
    loc = (rank % grid[0], (rank // grid[0]) % grid[1], rank // (grid[0] * grid[1]))
    nccl_id = lambda: None
    allsum = lambda a: a
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

        def nccl_id():
            """a fresh ncclUniqueId per context, made on rank 0 and broadcast (what MPI_Bcast does in a LAMMPS host)"""
            idt = torch.zeros(pkg.lib.NCCL_ID_BYTES, dtype=torch.uint8, device="cuda")
            if rank == 0:
                idt.copy_(torch.frombuffer(bytearray(pkg.lib.nccl_unique_id()), dtype=torch.uint8))
            dist.broadcast(idt, 0)
            return bytes(idt.cpu().numpy().tobytes())

        def allsum(a):
            t = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).cuda()
            dist.all_reduce(t)
            return t.cpu().numpy()

    have_ref = False
    if rank == 0 and not args.no_parity:
        import oracle  # the checker: only the parity / cpu_baseline legs touch it

        have_ref = oracle.ref_available()
    # ---- N > 1: decomposed parity on a small system first (the driver's pytest box has one GPU) ----
    parity_small = {}
    if world > 1 and not args.no_parity:
        for kind in ("lj",) if args.no_eam else ("lj", "eam"):
            try:
                r = run_decomposed_parity(torch, dist, pkg, kind, local, world, rank, nccl_id, grid, loc, allsum) if (have_ref or rank != 0) else None
            except Exception as e:  # noqa: BLE001
                r = {"error": repr(e), "ok": False}
            parity_small[kind] = r

    lj, md_lj = run_workload(torch, dist, pkg, "lj", args.cells, args, local, world, rank, nccl_id, grid, loc, allsum)
    eam, md_eam = None, None
    if not args.no_eam:
        eam, md_eam = run_workload(torch, dist, pkg, "eam", args.eam_cells, args, local, world, rank, nccl_id, grid, loc, allsum)

    chain = None
    if args.chain > 0 and world == 1:
        try:
            chain = run_chain(torch, dist, pkg, args.chain, args, local)
        except Exception as e:  # noqa: BLE001  (never lose the headline line over the third config)
            chain = {"error": repr(e)}

    # ---- parity at the quoted sizes: the reference's own thermo on the same input (rank 0, host cores) ----
    def lattice_check(w, kind):
        """size-independent step-0 facts of the perfect fcc lattice: potential energy per atom equals the 32 000-atom
        reference value (BASELINE.md, 15 digits) and the temperature is the `velocity create` target exactly"""
        pe_atom = {"lj": -6.77336805323422, "eam": -113280.000077752 / 32000.0}[kind]
        t_target = {"lj": 1.44, "eam": 1600.0}[kind]
        t0 = w["thermo0"]
        return {"pe_per_atom_rel": abs(t0["pe"] / w["natoms"] - pe_atom) / abs(pe_atom), "temp_rel": abs(t0["temp"] - t_target) / t_target,
                "natoms": w["natoms"]}

    parity = {}
    if rank == 0 and not args.no_parity:
        cores = os.cpu_count() or 1
        for kind, w, cells in (("lj", lj, args.cells), ("eam", eam, args.eam_cells)):
            if w is None:
                continue
            blk = {"lattice_step0": lattice_check(w, kind)}
            if world == 1 and have_ref and "parity_run" in w:
                try:
                    pr = w["parity_run"]
                    ref = ref_thermo(kind, (cells,) * 3, pr["steps"], cores)
                    blk.update(parity_block(f"oracle/_ref/lmp_ref (unmodified reference, OPENMP styles, {cores} threads) on the same "
                                            f"{cells}^3-cell input = {w['natoms']} atoms", w["thermo0"], pr["thermo"], pr["builds"], ref))
                except Exception as e:  # noqa: BLE001
                    blk.update({"error": repr(e), "ok": False})
            elif world > 1:
                small = parity_small.get(kind)
                if small is not None:
                    blk.update(small)
                blk["note"] = ("full size: lattice_step0 only (the reference cannot run the %d-atom box in bench time); decomposed parity "
                               "on the small system above" % w["natoms"])
            else:
                blk["note"] = "oracle/_ref/lmp_ref not available on this box"
            if "ok" in blk:
                blk["ok"] = bool(blk["ok"] and blk["lattice_step0"]["pe_per_atom_rel"] <= PARITY_TOL)
            parity[kind] = blk
        if chain is not None and "error" not in chain and have_ref:
            try:   # the thermostat generators differ after step 0 (counter-based on the device): setup state only
                ref = ref_thermo("chain", (args.chain,) * 3, 0, cores)
                t0 = dict(chain["thermo0"])
                parity["chain"] = parity_block(f"oracle/_ref/lmp_ref on bench/data.chain replicated {args.chain}^3, step 0 (setup forces: pair + "
                                               f"FENE; later steps differ by the thermostat's random stream)", t0, t0, 0, ref)
                parity["chain"]["builds"] = None
                parity["chain"]["ok"] = bool(max(parity["chain"]["step0"].values()) <= PARITY_TOL)
            except Exception as e:  # noqa: BLE001
                parity["chain"] = {"error": repr(e), "ok": False}

    def emit(e2e_error=None):
        if rank != 0:
            return
        peaks = {}
        try:
            peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if "hbm_gbs" in peaks else "B200_PROFILING.md fallback 6650 GB/s (of fallback)"

        try:  # measured DRAM bytes per atom of each kernel (one ncu --set full capture, committed under profiles/)
            dram = json.loads((ROOT / "profiles" / "r02_dram_traffic.json").read_text())
        except Exception:
            dram = {}

        def roof(w, cls, bytes_per_atom, dram_key=None, fused_nve=False):
            kms, kc = w["kernel_ms"][cls]
            if kc == 0 or kms <= 0.0:
                return None
            avg_ms = kms / kc
            rows = None
            if fused_nve:
                # fix nve runs INSIDE this kernel on the steps that neither tally nor end the run (INTEG variants): those
                # launches do two of SURVEY section 8d's rows.  Separate integrate launches seen in the timed steps: one
                # initial_integrate at the start plus one per unfused step.
                nve_launches = w["kernel_ms"]["nve"][1]
                fused = min(1.0, max(0.0, 1.0 - (nve_launches - 1) / max(args.steps, 1)))
                rows = {"pair_row_bytes_per_atom": bytes_per_atom, "nve_row_bytes_per_atom": BYTES_NVE, "fused_share_of_steps": fused,
                        "frac_pair_row_only": bytes_per_atom * w["atoms_per_gpu"] / (avg_ms * 1e-3) / 1e9 / peak}
                bytes_per_atom = bytes_per_atom + BYTES_NVE * fused
            ach = bytes_per_atom * w["atoms_per_gpu"] / (avg_ms * 1e-3) / 1e9
            r = {"bound": "hbm", "kernel": cls, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                 "avg_launch_ms": avg_ms, "launches": kc, "algorithmic_bytes_per_atom": bytes_per_atom, "peak_source": peak_src}
            if rows:
                r["rows"] = rows
            if dram_key in dram:
                r["traffic"] = dram[dram_key]["bytes_per_atom"] * w["atoms_per_gpu"]
                r["traffic_source"] = (f"ncu dram__bytes_read+write = {dram[dram_key]['bytes_per_atom']} B/atom on "
                                       f"{dram[dram_key]['capture_atoms']} atoms ({dram[dram_key].get('source', 'profiles/')}), scaled to this launch")
            return r

        line = {"metric": "Matom-steps/s (LJ melt, bench/in.lj scaled)", "value": lj["value"], "unit": "Matom-steps/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": lj["ms"] / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": workload_config(args.cells, world),
                "working_set_gb": lj["mem_gb"], "clocks": lj["clocks"], "gpu_launches": lj["launches"],
                "roofline": roof(lj, "pair", BYTES_LJ_PAIR, "pair_lj", fused_nve=True),
                "step_roofline": {"algorithmic_bytes_per_atom_step": BYTES_LJ_STEP,
                                  "frac_of_hbm_peak": BYTES_LJ_STEP * lj["atoms_per_gpu"] / (lj["ms"] / args.steps * 1e-3) / 1e9 / peak},
                "kernel_ms": lj["kernel_ms"], "ms_profiled_pass": lj["ms_profiled_pass"],
                "thermo": {"step0": lj["thermo0"], "last": lj["thermo1"]}, "neigh": lj["neigh"], "nghost": lj["nghost"], "halo": lj.get("halo")}
        line["e2e"] = lj.get("e2e")
        if parity:
            line["parity"] = parity.get("lj")
            line["parity_all_ok"] = bool(all(v.get("ok", True) for v in parity.values()))
        if e2e_error:
            line["e2e_error"] = e2e_error
        if eam is not None:
            line["eam"] = {"value": eam["value"], "unit": "Matom-steps/s", "ms_per_step": eam["ms"] / args.steps,
                           "workload": f"EAM Cu {eam['atoms_per_gpu']} atoms per GPU ({eam['natoms']} total; bench/in.eam scaled)",
                           "roofline": roof(eam, "pair", BYTES_EAM_FORCE, "pair_eam", fused_nve=True), "roofline_rho": roof(eam, "eam_rho", BYTES_EAM_RHO, "eam_rho"),
                           "step_frac_of_hbm_peak": BYTES_EAM_STEP * eam["atoms_per_gpu"] / (eam["ms"] / args.steps * 1e-3) / 1e9 / peak,
                           "kernel_ms": eam["kernel_ms"], "gpu_launches": eam["launches"], "clocks": eam["clocks"],
                           "thermo": {"step0": eam["thermo0"], "last": eam["thermo1"]}, "neigh": eam["neigh"], "e2e": eam.get("e2e"),
                           "parity": parity.get("eam")}
        if chain is not None:
            line["chain"] = chain if "error" in chain else {
                "value": chain["value"], "unit": "Matom-steps/s", "ms_per_step": chain["ms"] / args.steps,
                "workload": f"bench/in.chain replicated {args.chain}^3 = {chain['natoms']} beads on one GPU, counter-based thermostat",
                "kernel_ms": chain["kernel_ms"], "gpu_launches": chain["launches"], "neigh": chain["neigh"],
                "roofline": roof(chain, "neigh", BYTES_NEIGH_CHAIN), "roofline_pair": roof(chain, "pair+bond", BYTES_CHAIN_PAIR),
                "thermo": {"step0": chain["thermo0"], "last": chain["thermo1"]}, "parity": parity.get("chain")}
        if world == 1 and not args.no_e2e and not e2e_error and line["e2e"]:
            try:   # beside e2e: the loop time LAMMPS itself prints for the drop-in on the same input
                line["e2e"]["lammps_host"] = lammps_host_leg("lj", args.cells, args.steps, args.warmup)
            except Exception as ex:  # noqa: BLE001
                line["e2e"]["lammps_host"] = {"error": repr(ex)}
        if world == 1 and not args.no_cpu and not e2e_error:
            line["cpu_baseline"] = cpu_measure("lj", 100, 10, budget_s=20.0, cells_cap=args.cells)
            if eam is not None:
                line["eam"]["cpu_baseline"] = cpu_measure("eam", 100, 10, budget_s=15.0, cells_cap=args.eam_cells)
            if chain is not None and "error" not in chain:
                line["chain"]["cpu_baseline"] = cpu_measure("chain", 100, 10, budget_s=15.0, cells_cap=args.chain)

        def clean(o):
            if isinstance(o, dict):
                return {k: clean(v) for k, v in o.items()}
            if isinstance(o, (list, tuple)):
                return [clean(v) for v in o]
            if isinstance(o, np.ndarray):
                return o.tolist()
            if isinstance(o, (np.floating, np.integer)):
                return o.item()
            return o

        print(json.dumps(clean(line)), flush=True)

    # ---- end to end through host buffers, bounded: a failure here must not cost the numbers above ----
    if not args.no_e2e:
        with Watchdog(240.0, emit):
            try:
                run_e2e(torch, dist, md_lj, lj, args, local, world)
                if eam is not None:
                    run_e2e(torch, dist, md_eam, eam, args, local, world)
            except Exception as ex:  # the peers of a failed rank sit in a collective: say what we have and leave
                emit(f"{type(ex).__name__}: {ex}")
                sys.stdout.flush()
                os._exit(0)
    md_lj.close()
    if md_eam is not None:
        md_eam.close()
    emit()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
