"""Host-side mirror of the reference's style interface for the hot path, over the C ABI (include/b200md.h).

Same names, argument meaning and error behaviour as the reference's script commands / style classes:

    pair_style lj/cut Rc ; pair_coeff I J eps sigma [cut] ; pair_modify shift yes mix arithmetic
                                              -> PairLJCutB200.settings / coeff / modify   (src/pair_lj_cut.cpp:428-557)
    pair_style eam ; pair_coeff I I file      -> PairEAMB200.coeff                         (src/MANYBODY/pair_eam.cpp:372-412)
    pair_style eam/alloy | eam/fs ; pair_coeff * * file El1 El2 ...
                                              -> PairEAMAlloyB200.coeff                    (src/MANYBODY/pair_eam_alloy.cpp:55-110)
    neighbor SKIN bin ; neigh_modify every N delay M check yes|no
                                              -> NeighborB200.set / modify                 (src/neighbor.cpp:2556-2700)
    fix ID all nve                            -> FixNVEB200                                (src/fix_nve.cpp)
    bond_style fene ; bond_coeff ; special_bonds fene ; fix ID all langevin T T damp seed   (bench/in.chain)
                                              -> BondFENEB200, B200MD.bonds / fix("langevin", ...)
    timestep DT ; run N ; thermo N            -> B200MD.timestep / run                     (src/verlet.cpp:229-360)

The C++ versions of these classes for a LAMMPS build (suffix ``b200``) are described in INTEGRATION.md; this
module is what tests and bench.py drive.  All compute happens in libb200md.so; nothing here touches oracle/.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import lib
from .inputs import UNITS


class PairLJCutB200:
    """pair_style lj/cut (src/pair_lj_cut.cpp): settings(), coeff(), init_one() tables, handed to b200md_lj_init."""

    def __init__(self, owner: "B200MD"):
        self.o = owner
        n = owner.ntypes + 1
        self.cut_global = None
        self.setflag = np.zeros((n, n), dtype=bool)
        self.epsilon = np.zeros((n, n))
        self.sigma = np.zeros((n, n))
        self.cut = np.zeros((n, n))
        self.offset_flag = False
        self.mix_flag = "geometric"
        self.special_lj = np.array([1.0, 0.0, 0.0, 0.0])  # Force::special_lj default (src/force.cpp)

    def settings(self, cut_global: float):
        self.cut_global = float(cut_global)
        self.cut[self.setflag] = self.cut_global  # reset cutoffs that were set (src/pair_lj_cut.cpp:437-444)

    def coeff(self, i: int, j: int, eps: float, sigma: float, cut: float | None = None):
        if self.cut_global is None:
            raise ValueError("pair_coeff before pair_style settings")
        nt = self.o.ntypes
        if not (1 <= i <= nt and 1 <= j <= nt):
            raise ValueError("Incorrect args for pair coefficients")
        i, j = min(i, j), max(i, j)
        self.epsilon[i, j], self.sigma[i, j] = eps, sigma
        self.cut[i, j] = self.cut_global if cut is None else cut
        self.setflag[i, j] = True

    def modify(self, shift: bool | None = None, mix: str | None = None):
        if shift is not None:
            self.offset_flag = bool(shift)
        if mix is not None:
            if mix not in ("geometric", "arithmetic"):
                raise ValueError("Illegal pair_modify command")
            self.mix_flag = mix

    def _mix_distance(self, a, b):
        return math.sqrt(a * b) if self.mix_flag == "geometric" else 0.5 * (a + b)

    def init(self):
        """Pair::init + init_one for every i <= j (src/pair.cpp:215-301, src/pair_lj_cut.cpp:501-524)."""
        nt = self.o.ntypes
        n = nt + 1
        for i in range(1, n):
            if not self.setflag[i, i]:
                raise ValueError("All pair coeffs are not set")
        t = {k: np.zeros((n, n)) for k in ("cutsq", "lj1", "lj2", "lj3", "lj4", "offset")}
        for i in range(1, n):
            for j in range(i, n):
                if not self.setflag[i, j]:
                    self.epsilon[i, j] = math.sqrt(self.epsilon[i, i] * self.epsilon[j, j])
                    self.sigma[i, j] = self._mix_distance(self.sigma[i, i], self.sigma[j, j])
                    self.cut[i, j] = self._mix_distance(self.cut[i, i], self.cut[j, j])
                e, s, c = self.epsilon[i, j], self.sigma[i, j], self.cut[i, j]
                vals = dict(lj1=48.0 * e * math.pow(s, 12.0), lj2=24.0 * e * math.pow(s, 6.0), lj3=4.0 * e * math.pow(s, 12.0),
                            lj4=4.0 * e * math.pow(s, 6.0), cutsq=c * c, offset=0.0)
                if self.offset_flag and c > 0.0:
                    r = s / c
                    vals["offset"] = 4.0 * e * (math.pow(r, 12.0) - math.pow(r, 6.0))
                for k, v in vals.items():
                    t[k][i, j] = t[k][j, i] = v
                self.cut[j, i] = c
        self.tables = t
        self.o.ctx.call("b200md_lj_init", nt, *(lib.ptr(t[k]) for k in ("cutsq", "lj1", "lj2", "lj3", "lj4", "offset")),
                        lib.ptr(self.special_lj))
        return np.sqrt(t["cutsq"])  # per-pair force cutoffs, for Neighbor


class PairEAMB200:
    """pair_style eam, funcfl files (src/MANYBODY/pair_eam.cpp).  coeff(i, i, file) as the reference: only i == j,
    and it sets the mass of type i from the file."""

    def __init__(self, owner: "B200MD"):
        self.o = owner
        self.files: list[str] = []
        self.map = np.full(owner.ntypes + 1, -1, dtype=np.int32)
        self.cutmax = 0.0

    def settings(self):
        pass

    def coeff(self, i: int, j: int, filename: str):
        if i != j or not (1 <= i <= self.o.ntypes):
            raise ValueError("Incorrect args for pair coefficients")
        filename = str(filename)
        if filename not in self.files:
            self.files.append(filename)
        self.map[i] = self.files.index(filename)

    def init(self):
        nt = self.o.ntypes
        if (self.map[1:] < 0).all():
            raise ValueError("All pair coeffs are not set")
        arr = (C.c_char_p * len(self.files))(*[f.encode() for f in self.files])
        mass = np.zeros(nt + 1)
        cut = C.c_double(0.0)
        self.o.ctx.call("b200md_eam_init_funcfl", nt, len(self.files), arr, lib.ptr(self.map), lib.ptr(mass), C.byref(cut))
        for i in range(1, nt + 1):
            if self.map[i] >= 0:
                self.o.mass[i] = mass[i]
        self.cutmax = cut.value
        return np.full((nt + 1, nt + 1), self.cutmax)


class PairEAMAlloyB200:
    """pair_style eam/alloy (setfl) and eam/fs (src/MANYBODY/pair_eam_alloy.cpp, pair_eam_fs.cpp):
    coeff("*", "*", file, El1, El2, ...) with one element name (or "NULL") per atom type; sets the type masses."""

    def __init__(self, owner: "B200MD", fs: bool = False):
        self.o = owner
        self.fs = fs
        self.file = None
        self.elements: list[str] = []
        self.cutmax = 0.0

    def settings(self):
        pass

    def coeff(self, i, j, filename: str, *elements: str):
        if str(i) != "*" or str(j) != "*" or len(elements) != self.o.ntypes:
            raise ValueError("Incorrect args for pair coefficients")
        self.file = str(filename)
        self.elements = [str(e) for e in elements]

    def init(self):
        nt = self.o.ntypes
        if self.file is None:
            raise ValueError("All pair coeffs are not set")
        arr = (C.c_char_p * (nt + 1))(None, *[e.encode() for e in self.elements])
        mass = np.zeros(nt + 1)
        cut = C.c_double(0.0)
        self.o.ctx.call("b200md_eam_init_setfl", nt, self.file.encode(), int(self.fs), arr, lib.ptr(mass), C.byref(cut))
        for i in range(1, nt + 1):
            if self.elements[i - 1] != "NULL":
                self.o.mass[i] = mass[i]
        self.cutmax = cut.value
        return np.full((nt + 1, nt + 1), self.cutmax)


class NeighborB200:
    """neighbor / neigh_modify (src/neighbor.cpp:2556-2700); cutneighsq[i][j] = (cut_ij + skin)^2 (:330-344)."""

    def __init__(self, owner: "B200MD"):
        self.o = owner
        self.skin = UNITS[owner.units]["skin"]
        self.every, self.delay, self.check = 1, 0, 1  # defaults of this LAMMPS version (src/neighbor.cpp:93-98)
        self.binsize = 0.0

    def set(self, skin: float, style: str = "bin"):
        if style != "bin":
            raise ValueError("only `neighbor <skin> bin` is on the B200 path")
        if skin < 0:
            raise ValueError("Illegal neighbor command")
        self.skin = float(skin)

    def modify(self, every=None, delay=None, check=None, binsize=None):
        if every is not None:
            if every <= 0:
                raise ValueError("Illegal neigh_modify command")
            self.every = int(every)
        if delay is not None:
            if delay < 0:
                raise ValueError("Illegal neigh_modify command")
            self.delay = int(delay)
        if check is not None:
            self.check = 1 if check in (True, 1, "yes") else 0
        if binsize is not None:
            self.binsize = float(binsize)

    def init(self, cut: np.ndarray):
        nt = self.o.ntypes
        if self.delay > 0 and self.delay % self.every != 0:
            raise ValueError("Neighbor delay must be 0 or multiple of every setting")
        cn = np.zeros_like(cut)
        cn[1:, 1:] = cut[1:, 1:] + self.skin
        self.cutneighsq = cn * cn
        self.cutneighmax = float(cn[1:, 1:].max())
        self.o.ctx.call("b200md_neigh_config", nt, lib.ptr(self.cutneighsq), self.cutneighmax, self.skin, self.every,
                        self.delay, self.check, self.binsize)


class FixNVEB200:
    """fix nve (src/fix_nve.cpp): dtv = dt, dtf = 0.5*dt*ftm2v; step pieces exposed for the plugin-style loop."""

    def __init__(self, owner: "B200MD"):
        self.o = owner

    def init(self):
        self.dtv = self.o.dt
        self.dtf = 0.5 * self.o.dt * UNITS[self.o.units]["ftm2v"]

    def initial_integrate(self):
        self.o.ctx.call("b200md_nve_initial", self.dtv, self.dtf)

    def final_integrate(self):
        self.o.ctx.call("b200md_nve_final", self.dtf)


class BondFENEB200:
    """bond_style fene ; bond_coeff N K R0 epsilon sigma (src/MOLECULE/bond_fene.cpp:129-165)."""

    def __init__(self, owner: "B200MD", nbondtypes: int):
        self.o = owner
        self.n = int(nbondtypes)
        self.k = np.zeros(self.n + 1)
        self.r0 = np.zeros(self.n + 1)
        self.epsilon = np.zeros(self.n + 1)
        self.sigma = np.zeros(self.n + 1)
        self.setflag = np.zeros(self.n + 1, dtype=bool)

    def coeff(self, itype: int, k: float, r0: float, epsilon: float, sigma: float):
        if not 1 <= itype <= self.n:
            raise ValueError("Incorrect args for bond coefficients")
        self.k[itype], self.r0[itype], self.epsilon[itype], self.sigma[itype] = k, r0, epsilon, sigma
        self.setflag[itype] = True

    def init(self):
        if not self.setflag[1:].all():
            raise ValueError("All bond coeffs are not set")
        self.o.ctx.call("b200md_bond_fene", self.n, lib.ptr(self.k), lib.ptr(self.r0), lib.ptr(self.epsilon), lib.ptr(self.sigma))


class FixLangevinB200:
    """fix ID all langevin Tstart Tstop damp seed (src/fix_langevin.cpp:55-160), plain instance."""

    def __init__(self, owner: "B200MD", t_start: float, t_stop: float, damp: float, seed: int, rng: str = "mars"):
        if rng not in ("mars", "counter"):
            raise ValueError("fix langevin: rng is 'mars' (the reference's stream, one rank) or 'counter' (device, any ranks)")
        self.rng = rng
        if damp <= 0.0:
            raise ValueError("Fix langevin period must be > 0.0")
        if seed <= 0:
            raise ValueError("Illegal fix langevin command")
        self.o = owner
        self.t_start, self.t_stop, self.damp, self.seed = float(t_start), float(t_stop), float(damp), int(seed)

    def init(self):
        self.o.ctx.call("b200md_langevin_rng", 1 if self.rng == "counter" else 0)
        self.o.ctx.call("b200md_fix_langevin", self.t_start, self.t_stop, self.damp, self.seed, self.o.dt)


class B200MD:
    """The slice of a LAMMPS instance the hot path needs: box, atoms, one pair style, neighbor, fix nve, run."""

    def __init__(self, units: str = "lj", ntypes: int = 1, device: int = 0, nranks: int = 1, rank: int = 0,
                 nccl_id: bytes | None = None, procgrid=(1, 1, 1)):
        self.units = units
        self.ntypes = ntypes
        self.ctx = lib.Context(device, nranks, rank, nccl_id)
        self.nranks, self.rank = nranks, rank
        self.procgrid = np.asarray(procgrid, dtype=np.int32)
        u = UNITS[units]
        self.ctx.call("b200md_set_units", u["boltz"], u["mvv2e"], u["ftm2v"], u["nktv2p"])
        self.dt = u["dt"]
        self.mass = np.zeros(ntypes + 1)
        self.pair = None
        self.neighbor = NeighborB200(self)
        self.fix_nve = FixNVEB200(self)
        self.natoms_total = 0
        self._initialized = False
        self.thermo_log: list[dict] = []
        self.bond = None            # bond_style fene (bench/in.chain)
        self.fix_langevin = None
        self.sortfreq, self.sort_binsize = 1000, 0.0   # atom_modify sort defaults (src/atom.cpp:119-120)
        self._order = None
        self._special_lj = None

    # -- script-command analogues ---------------------------------------------------------------
    def box(self, lo, hi, periodic=(1, 1, 1)):
        self.boxlo = np.asarray(lo, dtype=np.float64)
        self.boxhi = np.asarray(hi, dtype=np.float64)
        self.periodic = np.asarray(periodic, dtype=np.int32)
        self.ctx.call("b200md_set_box", lib.ptr(self.boxlo), lib.ptr(self.boxhi), lib.ptr(self.periodic), lib.ptr(self.procgrid))

    def set_mass(self, itype: int, m: float):
        self.mass[itype] = m

    def create_atoms(self, x, v, type_, tag, image=None, natoms_total=None):
        x = np.ascontiguousarray(x, dtype=np.float64)
        v = None if v is None else np.ascontiguousarray(v, dtype=np.float64)
        type_ = np.ascontiguousarray(type_, dtype=np.int32)
        tag = np.ascontiguousarray(tag, dtype=np.int32)
        image = None if image is None else np.ascontiguousarray(image, dtype=np.int32)
        if self._order is None or self._order.shape[0] != tag.shape[0]:
            self._order = tag.copy()
        self.ctx.call("b200md_upload_atoms", x.shape[0], lib.ptr(x), lib.ptr(v), lib.ptr(type_), lib.ptr(tag), lib.ptr(image))
        self.natoms_total = int(natoms_total if natoms_total is not None else x.shape[0])
        self._initialized = False

    def bonds(self, tag, num_bond, bond_type, bond_atom, newton_bond: bool = True):
        """The molecular atom style's bond arrays in the caller's local order (read_data Bonds section ->
        Atom::data_bonds, src/atom.cpp:1380-1460)."""
        tag = np.ascontiguousarray(tag, dtype=np.int32)
        num_bond = np.ascontiguousarray(num_bond, dtype=np.int32)
        bond_type = np.ascontiguousarray(bond_type, dtype=np.int32).reshape(tag.shape[0], -1)
        bond_atom = np.ascontiguousarray(bond_atom, dtype=np.int32).reshape(tag.shape[0], -1)
        self.ctx.call("b200md_set_bonds", tag.shape[0], lib.ptr(tag), bond_type.shape[1], lib.ptr(num_bond), lib.ptr(bond_type),
                      lib.ptr(bond_atom), int(newton_bond))

    def bond_style(self, name: str, nbondtypes: int = 1):
        if name not in ("fene", "fene/b200"):
            raise ValueError(f"Unrecognized bond style '{name}' is not on the B200 hot path")
        self.bond = BondFENEB200(self, nbondtypes)
        self._initialized = False
        return self.bond

    def special_bonds(self, keyword: str):
        """Only `special_bonds fene` (lj 0.0 1.0 1.0): what the bonded neighbor build implements."""
        if keyword != "fene":
            raise ValueError("special_bonds: only the fene setting (lj 0.0 1.0 1.0) is on the B200 hot path")
        # Force::special_lj as `special_bonds fene` sets it (src/force.cpp: special_lj[1..3] = 0, 1, 1); the library
        # refuses a bonded system whose pair style was initialised with any other weights
        self._special_lj = np.array([1.0, 0.0, 1.0, 1.0])
        if getattr(self, "pair", None) is not None and hasattr(self.pair, "special_lj"):
            self.pair.special_lj[:] = self._special_lj
            self._initialized = False

    def fix(self, style: str, *args, **kw):
        if style in ("langevin", "langevin/b200"):
            self.fix_langevin = FixLangevinB200(self, *args, **kw)
            self._initialized = False
            return self.fix_langevin
        if style in ("nve", "nve/b200"):
            return self.fix_nve
        raise ValueError(f"Unrecognized fix style '{style}' is not on the B200 hot path")

    def atom_modify_sort(self, freq: int, binsize: float = 0.0):
        self.sortfreq, self.sort_binsize = int(freq), float(binsize)

    def local_order(self, tag):
        """atom->tag[0..nlocal) of the reference at this point of the script: the order fix langevin's random
        stream is consumed in (defaults to the order atoms were handed to create_atoms)."""
        self._order = np.ascontiguousarray(tag, dtype=np.int32)

    def pair_style(self, name: str, *args):
        if name in ("lj/cut", "lj/cut/b200"):
            self.pair = PairLJCutB200(self)
            self.pair.settings(*args)
            if getattr(self, "_special_lj", None) is not None:
                self.pair.special_lj[:] = self._special_lj
        elif name in ("eam", "eam/b200"):
            self.pair = PairEAMB200(self)
        elif name in ("eam/alloy", "eam/alloy/b200"):
            self.pair = PairEAMAlloyB200(self, fs=False)
        elif name in ("eam/fs", "eam/fs/b200"):
            self.pair = PairEAMAlloyB200(self, fs=True)
        else:
            raise ValueError(f"Unrecognized pair style '{name}' is not on the B200 hot path")
        self._initialized = False
        return self.pair

    def timestep(self, dt: float):
        self.dt = float(dt)

    def tune(self, key: str, value: int):
        """`package`-style execution knobs of the device context (include/b200md.h b200md_tune): overlap, host_pipeline, hs_min_chunk."""
        self.ctx.call("b200md_tune", key.encode(), int(value))

    # -- init / setup / run -------------------------------------------------------------------------
    def init(self):
        """LAMMPS::init order for this path: pair (tables, cutoffs) -> masses -> neighbor -> fix."""
        if self.pair is None:
            raise ValueError("No pair style defined")
        cut = self.pair.init()
        self.ctx.call("b200md_set_mass", self.ntypes, lib.ptr(self.mass))
        self.neighbor.init(cut)
        self.fix_nve.init()
        if self.bond is not None:
            self.bond.init()
        if self.fix_langevin is not None:
            self.fix_langevin.init()
            self.ctx.call("b200md_langevin_order", self._order.shape[0], lib.ptr(self._order), self.sortfreq, self.sort_binsize)
        self._initialized = True

    def setup(self):
        if not self._initialized:
            self.init()
        self.ctx.call("b200md_setup")
        self.thermo_log.append(self.thermo())

    def run(self, nsteps: int, thermo_every: int = 0):
        """`run N` with `thermo M`: device-resident Verlet loop; thermo recorded on multiples of M and on the last step."""
        if not self._initialized:
            self.init()
        if int(self.ctx.L.b200md_timestep(self.ctx.h)) == 0 and not self.thermo_log:
            self.setup()
        done = 0
        step0 = int(self.ctx.L.b200md_timestep(self.ctx.h))
        if self.fix_langevin is not None:   # update->beginstep / endstep of this `run` (src/run.cpp:172-176)
            self.ctx.call("b200md_run_bounds", step0, step0 + nsteps)
        while done < nsteps:
            step = int(self.ctx.L.b200md_timestep(self.ctx.h))
            chunk = nsteps - done
            if thermo_every > 0:
                chunk = min(chunk, thermo_every - step % thermo_every)
            self.ctx.call("b200md_run", chunk, self.dt, 0)
            done += chunk
            self.thermo_log.append(self.thermo())

    def thermo(self) -> dict:
        out = np.zeros(10)
        self.ctx.call("b200md_thermo", float(self.natoms_total), lib.ptr(out))
        th = dict(step=int(self.ctx.L.b200md_timestep(self.ctx.h)), pe=out[0], ke=out[1], temp=out[2], press=out[3],
                  virial=out[4:10].copy())
        if self.bond is not None:
            eb, vb = np.zeros(1), np.zeros(6)
            self.ctx.call("b200md_thermo_bond", lib.ptr(eb), lib.ptr(vb))
            th.update(evdwl=out[0], ebond=float(eb[0]), pe=out[0] + float(eb[0]), virial_bond=vb)
        return th

    # -- data access --------------------------------------------------------------------------------
    @property
    def nlocal(self):
        return int(self.ctx.L.b200md_nlocal(self.ctx.h))

    @property
    def nghost(self):
        return int(self.ctx.L.b200md_nghost(self.ctx.h))

    def gather(self, sort_by_tag: bool = True) -> dict:
        n = self.nlocal
        x, v, f = np.empty((n, 3)), np.empty((n, 3)), np.empty((n, 3))
        t, tag, img = np.empty(n, dtype=np.int32), np.empty(n, dtype=np.int32), np.empty(n, dtype=np.int32)
        self.ctx.call("b200md_download_atoms", lib.ptr(x), lib.ptr(v), lib.ptr(f), lib.ptr(t), lib.ptr(tag), lib.ptr(img))
        if sort_by_tag:
            o = np.argsort(tag, kind="stable")
            x, v, f, t, tag, img = x[o], v[o], f[o], t[o], tag[o], img[o]
        return dict(x=x, v=v, f=f, type=t, tag=tag, image=img)

    def halo_put(self, cap_records: int, allgather) -> None:
        """Switch the per-step halo to the fused pack-and-put kernel over peer memory (b200md_halo_put_export / _import +
        b200md_tune "halo_put"): `allgather(bytes) -> [bytes of rank 0, rank 1, ...]` is the host's collective
        (torch.distributed.all_gather_object, MPI_Allgather); cap_records = ghost records a landing buffer holds."""
        h = np.zeros(lib.IPC_HANDLE_BYTES, dtype=np.uint8)
        self.ctx.call("b200md_halo_put_export", int(cap_records), lib.ptr(h))
        blobs = allgather(h.tobytes() + np.int64(cap_records).tobytes())
        n = lib.IPC_HANDLE_BYTES
        handles = np.frombuffer(b"".join(b[:n] for b in blobs), dtype=np.uint8).copy()
        caps = np.array([np.frombuffer(b[n:n + 8], dtype=np.int64)[0] for b in blobs], dtype=np.int64)
        self.ctx.call("b200md_halo_put_import", len(blobs), lib.ptr(handles), lib.ptr(caps))
        self.tune("halo_put", 1)

    def gather_in_order(self, host_tag) -> dict:
        """owned atoms in the order of the caller's tag array (b200md_download_atoms_by_tag: the permutation runs on
        the device) -- what VerletB200 writes back into Atom::x / v / f / image"""
        host_tag = np.ascontiguousarray(host_tag, dtype=np.int32)
        n = len(host_tag)
        x, v, f, img = np.empty((n, 3)), np.empty((n, 3)), np.empty((n, 3)), np.empty(n, dtype=np.int32)
        self.ctx.call("b200md_download_atoms_by_tag", n, lib.ptr(host_tag), int(host_tag.max()), lib.ptr(x), lib.ptr(v), lib.ptr(f),
                      lib.ptr(img))
        return dict(x=x, v=v, f=f, tag=host_tag, image=img)

    def peratom(self, sort_by_tag: bool = True) -> dict:
        """compute pe/atom + compute stress/atom NULL pair: per-atom pair energy and virial (xx yy zz xy xz yz, energy
        units: stress/atom multiplies by -nktv2p) of the current configuration"""
        n = self.nlocal
        self.ctx.call("b200md_pair_peratom")
        e, v = np.empty(n), np.empty((n, 6))
        self.ctx.call("b200md_peratom_download", lib.ptr(e), lib.ptr(v))
        tag = self.gather(sort_by_tag=False)["tag"]
        if sort_by_tag:
            o = np.argsort(tag)
            e, v, tag = e[o], v[o], tag[o]
        return dict(eatom=e, vatom=v, tag=tag)

    def all_positions(self):
        n = self.nlocal + self.nghost
        x, tag = np.empty((n, 3)), np.empty(n, dtype=np.int32)
        self.ctx.call("b200md_download_all", lib.ptr(x), lib.ptr(tag))
        return x, tag

    def neighbor_list(self):
        """(numneigh[nlocal], neighbors[total]) in device order, as lammps_neighlist_element_neighbors exposes it."""
        n = self.nlocal
        numneigh = np.empty(n, dtype=np.int32)
        total = C.c_int64(0)
        self.ctx.call("b200md_neigh_export", lib.ptr(numneigh), None, C.byref(total))
        nb = np.empty(total.value, dtype=np.int32)
        self.ctx.call("b200md_neigh_export", lib.ptr(numneigh), lib.ptr(nb), C.byref(total))
        return numneigh, nb

    def neighbor_tag_pairs(self) -> np.ndarray:
        """The full list as directed (tag_i, tag_j) rows, one per stored entry."""
        numneigh, nb = self.neighbor_list()
        _, tag = self.all_positions()
        i = np.repeat(np.arange(len(numneigh)), numneigh)
        return np.stack([tag[i], tag[nb & 0x1FFFFFFF]], axis=1)

    def neigh_stats(self):
        nb, nd, mx = C.c_int64(0), C.c_int64(0), C.c_int(0)
        self.ctx.call("b200md_neigh_stats", C.byref(nb), C.byref(nd), C.byref(mx))
        return dict(builds=nb.value, dangerous=nd.value, maxneigh=mx.value)

    def timers(self, reset=False):
        ms = np.zeros(5)
        k = C.c_int64(0)
        self.ctx.call("b200md_timers", lib.ptr(ms), C.byref(k), int(reset))
        return dict(pair=ms[0], neigh=ms[1], comm=ms[2], modify=ms[3], other=ms[4], launches=k.value)

    def close(self):
        self.ctx.close()
