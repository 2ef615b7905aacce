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
    ap.add_argument("--chain", type=int, default=0, metavar="R",
                    help="also run bench/in.chain replicated R x R x R (32 000 R^3 beads; one GPU only; opt-in: the bonded "
                         "path has not run on hardware yet)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
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
    stream, barrier + synchronize on both sides, max over ranks.  Returns ms for the K steps."""
    ext = torch.cuda.ExternalStream(md.ctx.L.b200md_stream(md.ctx.h), device=device)
    md.ctx.call("b200md_run", warmup, md.dt, 0)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(device)
    if nranks > 1:
        dist.barrier()
    with torch.cuda.stream(ext):
        e0.record()
    md.ctx.call("b200md_run", steps, md.dt, 0)
    with torch.cuda.stream(ext):
        e1.record()
    torch.cuda.synchronize(device)
    if nranks > 1:
        dist.barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=device)
    if nranks > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


def time_host_buffers(torch, dist, md, steps, warmup, device, nranks):
    """The same steps through the host-buffer entry point (b200md_step_host): every step copies x, v from pinned
    host arrays to the device and x, v, f back -- what a Pair/Fix plugin under the stock host-resident Verlet
    loop pays (the reference GPU package's model).  Returns (ms for K steps, h2d bytes/step, d2h bytes/step)."""
    n = md.nlocal
    cap = n + n // 64 + 1024  # room for atoms that migrate in at a rebuild (several ranks)
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
    return float(ms[0].item()), int(ms[1].item()), int(ms[2].item())


def run_workload(torch, dist, pkg, kind, cells, args, device, nranks, rank, nccl_id, grid, loc, allsum):
    t_gen = time.perf_counter()
    x, v, tag, box, natoms_total = make_atoms(pkg, kind, cells, grid, loc, allsum)
    t_gen = time.perf_counter() - t_gen
    md = make_system(pkg, kind, device, nranks, rank, nccl_id(), grid)
    md.box([0, 0, 0], box)
    md.create_atoms(x, v, np.ones(len(x), dtype=np.int32), tag, natoms_total=natoms_total)
    del x, v, tag
    md.setup()
    out = dict(natoms=natoms_total, atoms_per_gpu=md.nlocal, thermo0=md.thermo_log[-1])

    # ---- timed region: device-resident steps ---------------------------------------------------------
    md.timers(reset=True)
    clocks = Clocks(device)
    if rank == 0:
        clocks.start()
    ms = time_resident(torch, dist, md, args.steps, args.warmup, device, nranks)
    out["clocks"] = clocks.stop() if rank == 0 else None
    launches_total = md.timers()["launches"]
    out["ms"] = ms
    out["launches"] = int(round(launches_total * args.steps / (args.steps + args.warmup)))
    out["value"] = natoms_total * args.steps / (ms * 1e-3) / 1e6

    # ---- second pass with per-kernel CUDA events (no host syncs added) for the roofline ---------------
    md.ctx.call("b200md_profile", 1)
    kernel_times(md, reset=True)
    ms2 = time_resident(torch, dist, md, args.steps, 0, device, nranks)
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
    out = dict(natoms=natoms, thermo0=md.thermo_log[-1])
    md.timers(reset=True)
    ms = time_resident(torch, dist, md, args.steps, args.warmup, device, 1)
    out["ms"] = ms
    out["value"] = natoms * args.steps / (ms * 1e-3) / 1e6
    out["launches"] = int(round(md.timers()["launches"] * args.steps / (args.steps + args.warmup)))
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


def run_e2e(torch, dist, md, out, args, device, nranks):
    """End to end through host buffers: every rank moves its own atoms over its own PCIe link every step."""
    k = max(3, min(args.steps, 20))
    ems, bi, bo = time_host_buffers(torch, dist, md, k, 3, device, nranks)
    out["e2e"] = dict(value=out["natoms"] * k / (ems * 1e-3) / 1e6, unit="Matom-steps/s", h2d_bytes_per_step=bi,
                      d2h_bytes_per_step=bo, steps=k, api="b200md_step_host (pinned host x,v in; x,v,f out every step)",
                      form="three-stream pipeline")


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
REF_IN = {
    "lj": """units lj
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
thermo 0
run {w}
run {k}
""",
    "eam": """units metal
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
thermo 0
run {w}
run {k}
""",
}


def ref_run(kind, n, steps, warmup, threads):
    """One launch of the unmodified reference on an n^3-cell sample: `run W` then `run K`; returns
    (Matom-steps/s of the K-step loop as Finish prints it, loop seconds)."""
    import oracle

    with tempfile.TemporaryDirectory() as tmp:
        inp = Path(tmp) / "in.bench"
        inp.write_text(REF_IN[kind].format(n=n, w=warmup, k=steps, pot=POT))
        cmd = [str(oracle.LMP_REF), "-in", str(inp), "-log", "none", "-echo", "none"]
        if threads > 1:
            cmd += ["-sf", "omp", "-pk", "omp", str(threads)]
        env = dict(os.environ, OMP_NUM_THREADS=str(threads), OMP_PROC_BIND="true")
        txt = subprocess.run(cmd, cwd=tmp, env=env, capture_output=True, text=True, check=True).stdout
    loops = re.findall(r"Loop time of ([0-9.eE+-]+) on \d+ procs for (\d+) steps with (\d+) atoms", txt)
    sec, st, na = loops[-1]
    return int(na) * int(st) / float(sec) / 1e6, float(sec)


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
    """Size an n^3-cell sample of the workload so that W+K steps take about `budget_s` on this host, run it."""
    import oracle

    cores = os.cpu_count() or 1
    have_ref = oracle.ref_available()
    if have_ref:
        rate, _ = ref_run(kind, 20, 10, 2, cores)  # calibration: the shipped 32k case, 10 steps
        kindname = "reference"
    else:
        oracle.build()
        rate, _ = port_run(kind, 10, 5, 1)
        cores, kindname = 1, "port"
    atoms = rate * 1e6 * budget_s / max(1, steps + warmup)
    n = int(max(10, min(cells_cap, round((atoms / 4) ** (1 / 3)))))
    val, sec = ref_run(kind, n, steps, warmup, cores) if have_ref else port_run(kind, n, steps, warmup)
    sample = (f"{'bench/in.' + kind} scaled to {n}^3 fcc cells = {4 * n ** 3} atoms, {steps} steps after {warmup} warm-up, "
              + (f"oracle/_ref/lmp_ref -sf omp -pk omp {cores}" if have_ref else "oracle/oracle.c port, 1 thread")
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
            "config": {"workload": f"LJ melt 16.384 M atoms per GPU (bench/in.lj -var x 8 -var y 8 -var z 8), rc=2.5, skin 0.3, "
                                   f"rebuild every 20; CPU arm timed on a bounded sample: {cb['sample']}"},
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

    lj, md_lj = run_workload(torch, dist, pkg, "lj", args.cells, args, local, world, rank, nccl_id, grid, loc, allsum)
    eam, md_eam = None, None
    if not args.no_eam:
        eam, md_eam = run_workload(torch, dist, pkg, "eam", args.eam_cells, args, local, world, rank, nccl_id, grid, loc, allsum)

    chain = None
    if args.chain > 0 and world == 1:
        try:
            chain = run_chain(torch, dist, pkg, args.chain, args, local)
        except Exception as e:  # noqa: BLE001  (opt-in extra: never lose the headline line over it)
            chain = {"error": repr(e)}

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
            dram = json.loads((ROOT / "profiles" / "r01_dram_traffic.json").read_text())
        except Exception:
            dram = {}

        def roof(w, cls, bytes_per_atom, dram_key=None):
            kms, kc = w["kernel_ms"][cls]
            if kc == 0 or kms <= 0.0:
                return None
            avg_ms = kms / kc
            ach = bytes_per_atom * w["atoms_per_gpu"] / (avg_ms * 1e-3) / 1e9
            r = {"bound": "hbm", "kernel": cls, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak, "traffic": None,
                 "avg_launch_ms": avg_ms, "launches": kc, "algorithmic_bytes_per_atom": bytes_per_atom, "peak_source": peak_src}
            if dram_key in dram:
                r["traffic"] = dram[dram_key]["bytes_per_atom"] * w["atoms_per_gpu"]
                r["traffic_source"] = (f"ncu dram__bytes_read+write = {dram[dram_key]['bytes_per_atom']} B/atom on "
                                       f"{dram[dram_key]['capture_atoms']} atoms (profiles/r01_ncu_pair_kernels.md), scaled to this launch")
            return r

        line = {"metric": "Matom-steps/s (LJ melt, bench/in.lj scaled)", "value": lj["value"], "unit": "Matom-steps/s", "n_gpus": world,
                "steps": args.steps, "warmup": args.warmup, "ms_per_step": lj["ms"] / args.steps, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                "config": {"workload": f"LJ melt {lj['atoms_per_gpu']} atoms per GPU ({lj['natoms']} total; bench/in.lj scaled, fcc 0.8442, "
                                       f"T=1.44 loop geom, lj/cut 2.5, skin 0.3, rebuild every 20 check no, fix nve, dt 0.005)",
                           "grid": list(grid), "l2": "inputs larger than L2 (per-GPU working set %.1f GB)" % lj["mem_gb"]},
                "clocks": lj["clocks"], "gpu_launches": lj["launches"],
                "roofline": roof(lj, "pair", BYTES_LJ_PAIR, "pair_lj"),
                "step_roofline": {"algorithmic_bytes_per_atom_step": BYTES_LJ_STEP,
                                  "frac_of_hbm_peak": BYTES_LJ_STEP * lj["atoms_per_gpu"] / (lj["ms"] / args.steps * 1e-3) / 1e9 / peak},
                "kernel_ms": lj["kernel_ms"], "ms_profiled_pass": lj["ms_profiled_pass"],
                "thermo": {"step0": lj["thermo0"], "last": lj["thermo1"]}, "neigh": lj["neigh"], "nghost": lj["nghost"]}
        line["e2e"] = lj.get("e2e")
        if e2e_error:
            line["e2e_error"] = e2e_error
        if eam is not None:
            line["eam"] = {"value": eam["value"], "unit": "Matom-steps/s", "ms_per_step": eam["ms"] / args.steps,
                           "workload": f"EAM Cu {eam['atoms_per_gpu']} atoms per GPU ({eam['natoms']} total; bench/in.eam scaled)",
                           "roofline": roof(eam, "pair", BYTES_EAM_FORCE, "pair_eam"), "roofline_rho": roof(eam, "eam_rho", BYTES_EAM_RHO, "eam_rho"),
                           "step_frac_of_hbm_peak": BYTES_EAM_STEP * eam["atoms_per_gpu"] / (eam["ms"] / args.steps * 1e-3) / 1e9 / peak,
                           "kernel_ms": eam["kernel_ms"], "gpu_launches": eam["launches"], "clocks": eam["clocks"],
                           "thermo": {"step0": eam["thermo0"], "last": eam["thermo1"]}, "neigh": eam["neigh"], "e2e": eam.get("e2e")}
        if chain is not None:
            line["chain"] = chain if "error" in chain else {
                "value": chain["value"], "unit": "Matom-steps/s", "ms_per_step": chain["ms"] / args.steps,
                "workload": f"bench/in.chain replicated {args.chain}^3 = {chain['natoms']} beads on one GPU, counter-based thermostat",
                "kernel_ms": chain["kernel_ms"], "gpu_launches": chain["launches"], "neigh": chain["neigh"],
                "thermo": {"step0": chain["thermo0"], "last": chain["thermo1"]}}
        if world == 1 and not args.no_cpu and not e2e_error:
            line["cpu_baseline"] = cpu_measure("lj", 100, 10, budget_s=20.0, cells_cap=args.cells)
            if eam is not None:
                line["eam"]["cpu_baseline"] = cpu_measure("eam", 100, 10, budget_s=15.0, cells_cap=args.eam_cells)

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
