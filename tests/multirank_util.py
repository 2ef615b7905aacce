"""Helpers that run the decomposed (one rank per GPU) path in spawned processes and bring results back."""
from __future__ import annotations

import importlib
import os
import sys
import tempfile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))


def _worker(rank, world, grid, kind, cells, steps, port, outdir, backend):
    import torch
    import torch.distributed as dist

    import bench

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group(backend, rank=rank, world_size=world)
    pkg = importlib.import_module("lammps-fugaku-threadpool_b200")
    loc = (rank % grid[0], (rank // grid[0]) % grid[1], rank // (grid[0] * grid[1]))

    def allsum(a):
        t = torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64))
        if backend == "nccl":
            t = t.cuda(rank)
        dist.all_reduce(t)
        return t.cpu().numpy()

    x, v, tag, box, natoms = bench.make_atoms(pkg, kind, cells, grid, loc, allsum)
    out = dict(x0=x, v0=v, tag0=tag, box=np.array(box), natoms=np.array([natoms]))
    if backend == "nccl":
        torch.cuda.set_device(rank)
        ids = [pkg.lib.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, 0)
        md = bench.make_system(pkg, kind, rank, world, rank, ids[0], grid)
        md.box([0, 0, 0], box)
        md.create_atoms(x, v, np.ones(len(x), dtype=np.int32), tag, natoms_total=natoms)
        if os.environ.get("B200MD_TEST_HALO_PUT"):   # per-step halo as the fused pack-and-put kernel over peer memory
            def allgather(blob):
                got = [None] * world
                dist.all_gather_object(got, blob)
                return got
            md.halo_put(int(os.environ["B200MD_TEST_HALO_PUT"]), allgather)
        md.setup()
        t0 = md.thermo_log[-1]
        numneigh, _ = md.neighbor_list()
        out.update(nneigh0=np.array([int(numneigh.sum())]), nghost0=np.array([md.nghost]))
        md.run(steps)
        t1 = md.thermo_log[-1]
        g = md.gather()
        st = md.neigh_stats()
        out.update(thermo0=np.array([t0["pe"], t0["ke"], t0["press"]]), thermo1=np.array([t1["pe"], t1["ke"], t1["press"]]),
                   x1=g["x"], v1=g["v"], f1=g["f"], tag1=g["tag"], builds=np.array([st["builds"], st["dangerous"]]),
                   nlocal1=np.array([md.nlocal]), puts=np.array([int(md.ctx.L.b200md_halo_put_count(md.ctx.h))]))
        md.close()
    np.savez(Path(outdir) / f"rank{rank}.npz", **out)
    dist.barrier()
    dist.destroy_process_group()


def run_decomposed(world, grid, kind, cells, steps, backend="nccl", port=29611):
    """Spawn `world` ranks; returns the list of per-rank result dicts."""
    import torch.multiprocessing as mp

    with tempfile.TemporaryDirectory() as tmp:
        mp.spawn(_worker, args=(world, tuple(grid), kind, cells, steps, port, tmp, backend), nprocs=world, join=True)
        res = []
        for r in range(world):
            d = np.load(Path(tmp) / f"rank{r}.npz")
            res.append({k: d[k] for k in d.files})
    return res
