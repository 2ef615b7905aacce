"""TEST INFRASTRUCTURE ONLY -- ctypes view of oracle/oracle.c (the plain-C CPU restatement of the
reference's short-range MD path) and of the unmodified reference built into oracle/_ref/.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` leg may import
this module.  The product package never does (tests/test_no_oracle_in_product.py greps for it).
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent
LIB = HERE / "liboracle.so"
REF_DIR = HERE / "_ref"
LMP_REF = REF_DIR / "lmp_ref"
REF_EXTRACT = REF_DIR / "ref_extract"
REFERENCE_ROOT = Path("/root/reference")


def build(force: bool = False) -> Path:
    """Compile oracle.c -> liboracle.so (no FMA contraction, like the reference's x86-64 -O3 build)."""
    src = HERE / "oracle.c"
    if force or not LIB.exists() or LIB.stat().st_mtime < src.stat().st_mtime:
        subprocess.check_call(
            ["gcc", "-O2", "-fPIC", "-shared", "-ffp-contract=off", "-std=c99", "-D_GNU_SOURCE",
             str(src), "-o", str(LIB), "-lm"])
    return LIB


def build_ref(jobs: int = 8) -> bool:
    """Compile the unmodified reference into oracle/_ref (only where /root/reference exists)."""
    if LMP_REF.exists() and REF_EXTRACT.exists():
        return True
    if not (REFERENCE_ROOT / "src").is_dir():
        return False
    subprocess.check_call(["make", "-f", "oracle/Makefile.ref", f"-j{jobs}"], cwd=str(ROOT),
                          stdout=subprocess.DEVNULL)
    return LMP_REF.exists()


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(str(LIB))
        dp, ip, vp = C.POINTER(C.c_double), C.POINTER(C.c_int), C.c_void_p
        L.orc_lattice_scale_lj.restype = C.c_double
        L.orc_lattice_scale_lj.argtypes = [C.c_double]
        L.orc_create_fcc.restype = C.c_int
        L.orc_create_fcc.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, vp, vp]
        L.orc_velocity_create_geom.argtypes = [C.c_int, vp, vp, vp, C.c_double, C.c_int, C.c_double, C.c_double, vp]
        L.orc_create.restype = vp
        L.orc_create.argtypes = [C.c_int]
        L.orc_destroy.argtypes = [vp]
        L.orc_set_units.argtypes = [vp, C.c_char_p]
        L.orc_set_dt.argtypes = [vp, C.c_double]
        L.orc_set_box.argtypes = [vp, vp, vp]
        L.orc_set_mass.argtypes = [vp, C.c_int, C.c_double]
        L.orc_set_neighbor.argtypes = [vp, C.c_double, C.c_int, C.c_int, C.c_int]
        L.orc_set_atoms.argtypes = [vp, C.c_int, vp, vp, vp, vp]
        L.orc_set_special.argtypes = [vp, C.c_int, vp, vp, vp]
        L.orc_pair_lj_cut.argtypes = [vp, vp, vp, vp, C.c_int]
        L.orc_eam_read_funcfl.restype = C.c_int
        L.orc_eam_read_funcfl.argtypes = [vp, C.c_char_p]
        L.orc_eam_map_type.argtypes = [vp, C.c_int, C.c_int]
        L.orc_pair_eam_init.argtypes = [vp]
        L.orc_eam_dims.argtypes = [vp, vp]
        L.orc_eam_scalars.argtypes = [vp, vp]
        L.orc_eam_tables.argtypes = [vp] + [vp] * 6
        L.orc_get_mass.restype = C.c_double
        L.orc_get_mass.argtypes = [vp, C.c_int]
        for name in ("orc_pbc", "orc_borders", "orc_forward_comm", "orc_reverse_comm", "orc_setup_bins",
                     "orc_bin_atoms", "orc_neigh_build", "orc_neighbor_build", "orc_nve_initial", "orc_nve_final",
                     "orc_setup"):
            getattr(L, name).argtypes = [vp]
        L.orc_coord2bin.restype = C.c_int
        L.orc_coord2bin.argtypes = [vp, vp]
        L.orc_bin_info.argtypes = [vp, vp]
        L.orc_neighbor_decide.restype = C.c_int
        L.orc_neighbor_decide.argtypes = [vp]
        L.orc_compute_forces.argtypes = [vp, C.c_int, C.c_int]
        L.orc_run.argtypes = [vp, C.c_int, C.c_int, C.c_int]
        L.orc_thermo.argtypes = [vp, vp]
        for name in ("orc_nlocal", "orc_nghost", "orc_ncalls", "orc_ndanger"):
            getattr(L, name).restype = C.c_int
            getattr(L, name).argtypes = [vp]
        L.orc_npairs.restype = C.c_int64
        L.orc_npairs.argtypes = [vp]
        L.orc_get_atoms.argtypes = [vp] + [vp] * 6
        L.orc_get_ghosts.argtypes = [vp, vp, vp]
        L.orc_get_pairs.argtypes = [vp, vp]
        # in.chain additions
        L.orc_enable_sort.argtypes = [vp, C.c_int]
        L.orc_sort_atoms.argtypes = [vp]
        L.orc_set_bonds.argtypes = [vp, C.c_int, vp, vp, vp]
        L.orc_bond_fene.argtypes = [vp, C.c_int, vp, vp, vp, vp]
        L.orc_fix_langevin.argtypes = [vp, C.c_double, C.c_double, C.c_double, C.c_int]
        L.orc_thermo_bond.argtypes = [vp, vp]
        L.orc_nbondlist.restype = C.c_int
        L.orc_nbondlist.argtypes = [vp]
        L.orc_ranmars.argtypes = [C.c_int, C.c_int, vp]
        L.orc_pair_eam_setfl.argtypes = [vp, C.c_char_p, C.c_int, vp]
        _lib = L
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


UNITS = {
    # src/update.cpp:145-207
    "lj": dict(boltz=1.0, mvv2e=1.0, ftm2v=1.0, nktv2p=1.0, dt=0.005, skin=0.3),
    "metal": dict(boltz=8.617343e-5, mvv2e=1.0364269e-4, ftm2v=1.0 / 1.0364269e-4, nktv2p=1.6021765e6, dt=0.001,
                  skin=2.0),
    "real": dict(boltz=0.0019872067, mvv2e=48.88821291 * 48.88821291, ftm2v=1.0 / 48.88821291 / 48.88821291,
                 nktv2p=68568.415, dt=1.0, skin=2.0),
}


def lattice_scale_lj(rhostar: float) -> float:
    return lib().orc_lattice_scale_lj(rhostar)


def create_fcc(nx: int, ny: int, nz: int, scale: float):
    n = 4 * nx * ny * nz
    x = np.empty((n, 3), dtype=np.float64)
    tag = np.empty(n, dtype=np.int32)
    m = lib().orc_create_fcc(nx, ny, nz, scale, _p(x), _p(tag))
    assert m == n, (m, n)
    return x, tag


def velocity_create_geom(x, type_, mass, t_desired, seed, units="lj"):
    u = UNITS[units]
    n = x.shape[0]
    v = np.empty((n, 3), dtype=np.float64)
    mass = np.ascontiguousarray(mass, dtype=np.float64)
    type_ = np.ascontiguousarray(type_, dtype=np.int32)
    lib().orc_velocity_create_geom(n, _p(np.ascontiguousarray(x)), _p(type_), _p(mass), t_desired, seed, u["mvv2e"],
                                   u["boltz"], _p(v))
    return v


class OracleSystem:
    """One rank, orthogonal periodic box, newton on -- the reference's default CPU path."""

    def __init__(self, ntypes: int, units: str = "lj"):
        self.L = lib()
        self.h = self.L.orc_create(ntypes)
        self.ntypes = ntypes
        self.units = units
        self.L.orc_set_units(self.h, units.encode())
        self.step = 0

    def __del__(self):
        try:
            if self.h:
                self.L.orc_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def set_box(self, lo, hi):
        lo = np.asarray(lo, dtype=np.float64)
        hi = np.asarray(hi, dtype=np.float64)
        self.L.orc_set_box(self.h, _p(lo), _p(hi))

    def set_dt(self, dt):
        self.L.orc_set_dt(self.h, dt)

    def set_mass(self, itype, m):
        self.L.orc_set_mass(self.h, itype, m)

    def mass(self, itype):
        return self.L.orc_get_mass(self.h, itype)

    def set_neighbor(self, skin, every=1, delay=0, check=1):
        """Must be called BEFORE the pair style is defined (cutneigh = cut + skin is fixed there)."""
        self.L.orc_set_neighbor(self.h, skin, every, delay, check)

    def set_atoms(self, x, v, type_, tag):
        x = np.ascontiguousarray(x, dtype=np.float64)
        v = None if v is None else np.ascontiguousarray(v, dtype=np.float64)
        type_ = np.ascontiguousarray(type_, dtype=np.int32)
        tag = np.ascontiguousarray(tag, dtype=np.int32)
        self.L.orc_set_atoms(self.h, x.shape[0], _p(x), _p(v), _p(type_), _p(tag))

    def set_special(self, nspecial, special, special_lj):
        nspecial = np.ascontiguousarray(nspecial, dtype=np.int32)
        special = np.ascontiguousarray(special, dtype=np.int32)
        sl = np.ascontiguousarray(special_lj, dtype=np.float64)
        self.L.orc_set_special(self.h, special.shape[1], _p(nspecial), _p(special), _p(sl))

    # ---- in.chain additions (atom_style bond, special_bonds fene, bond_style fene, fix langevin) ----
    def enable_sort(self, flag=True):
        """Atom::sort at setup, as the reference does (fixes the local order fix langevin draws its numbers in)."""
        self.L.orc_enable_sort(self.h, int(flag))

    def set_bonds(self, bonds):
        """bonds: (nbonds, 3) int rows `type a1 a2` as in a data file; stored on a1 (newton_bond on).  Also sets the
        1-2 special partners both ways; with `special_bonds fene` (lj 0 1 1) they are excluded from the pair list."""
        bonds = np.asarray(bonds, dtype=np.int64)
        a = self.atoms()
        n = self.nlocal
        where = {int(t): i for i, t in enumerate(a["tag"])}
        per = [[] for _ in range(n)]
        spec = [[] for _ in range(n)]
        for t, a1, a2 in bonds:
            per[where[int(a1)]].append((int(t), int(a2)))
            spec[where[int(a1)]].append(int(a2))
            spec[where[int(a2)]].append(int(a1))
        bpa = max(1, max(len(p) for p in per))
        num = np.array([len(p) for p in per], dtype=np.int32)
        btype = np.zeros((n, bpa), dtype=np.int32)
        batom = np.zeros((n, bpa), dtype=np.int32)
        for i, p in enumerate(per):
            for m, (t, a2) in enumerate(p):
                btype[i, m], batom[i, m] = t, a2
        self.L.orc_set_bonds(self.h, bpa, _p(num), _p(btype), _p(batom))
        ms = max(1, max(len(p) for p in spec))
        nspecial = np.zeros((n, 3), dtype=np.int32)
        special = np.zeros((n, ms), dtype=np.int32)
        for i, p in enumerate(spec):
            nspecial[i, :] = len(p)  # Special::build stops after 1-2 when the 1-3/1-4 weights are 1.0 (src/special.cpp)
            special[i, :len(p)] = p
        self.set_special(nspecial, special, [1.0, 0.0, 1.0, 1.0])

    def bond_fene(self, coeffs):
        """coeffs: {bondtype: (K, R0, epsilon, sigma)} like `bond_coeff`"""
        nt = max(coeffs)
        arr = np.zeros((4, nt + 1))
        for t, c in coeffs.items():
            arr[:, t] = c
        k, r0, e, sg = (np.ascontiguousarray(arr[q]) for q in range(4))
        self.L.orc_bond_fene(self.h, nt, _p(k), _p(r0), _p(e), _p(sg))

    def fix_langevin(self, t_start, t_stop, damp, seed):
        self.L.orc_fix_langevin(self.h, t_start, t_stop, damp, seed)

    def thermo_bond(self):
        out = np.zeros(7)
        self.L.orc_thermo_bond(self.h, _p(out))
        return dict(e_bond=out[0], virial_bond=out[1:].copy())

    def pair_lj_cut(self, epsilon, sigma, cut, offset_flag=False):
        """epsilon/sigma/cut: (ntypes+1)x(ntypes+1) tables (row/col 0 unused), already mixed."""
        e = np.ascontiguousarray(epsilon, dtype=np.float64)
        s = np.ascontiguousarray(sigma, dtype=np.float64)
        c = np.ascontiguousarray(cut, dtype=np.float64)
        self.L.orc_pair_lj_cut(self.h, _p(e), _p(s), _p(c), int(offset_flag))

    def pair_eam(self, files_by_type):
        """files_by_type: {itype: funcfl filename}; like one `pair_coeff i i file` per type."""
        loaded = {}
        for itype, fn in sorted(files_by_type.items()):
            if fn not in loaded:
                k = self.L.orc_eam_read_funcfl(self.h, str(fn).encode())
                if k < 0:
                    raise OSError(f"cannot read funcfl file {fn} ({k})")
                loaded[fn] = k
            self.L.orc_eam_map_type(self.h, itype, loaded[fn])
        self.L.orc_pair_eam_init(self.h)

    def pair_eam_setfl(self, filename, elements, fs=False):
        """pair_style eam/alloy (fs=False) or eam/fs; pair_coeff * * filename elements[0] elements[1] ... (one name per
        atom type, "NULL" allowed)"""
        assert len(elements) == self.ntypes
        arr = (C.c_char_p * (self.ntypes + 1))(None, *[e.encode() for e in elements])
        rc = self.L.orc_pair_eam_setfl(self.h, str(filename).encode(), int(fs), arr)
        if rc != 0:
            raise OSError(f"cannot use setfl/fs file {filename} ({rc})")

    def eam_tables(self):
        dims = np.zeros(5, dtype=np.int32)
        self.L.orc_eam_dims(self.h, _p(dims))
        nrho, nr, nfrho, nrhor, nz2r = (int(v) for v in dims)
        sc = np.zeros(5)
        self.L.orc_eam_scalars(self.h, _p(sc))
        ld = self.ntypes + 1
        frho = np.zeros((nfrho, nrho + 1, 7))
        rhor = np.zeros((nrhor, nr + 1, 7))
        z2r = np.zeros((nz2r, nr + 1, 7))
        t2f = np.zeros(ld, dtype=np.int32)
        t2r = np.zeros((ld, ld), dtype=np.int32)
        t2z = np.zeros((ld, ld), dtype=np.int32)
        self.L.orc_eam_tables(self.h, _p(frho), _p(rhor), _p(z2r), _p(t2f), _p(t2r), _p(t2z))
        return dict(nrho=nrho, nr=nr, rdr=sc[0], rdrho=sc[1], rhomax=sc[2], cutforcesq=sc[3], cutmax=sc[4],
                    frho_spline=frho, rhor_spline=rhor, z2r_spline=z2r, type2frho=t2f, type2rhor=t2r, type2z2r=t2z)

    # ---- pieces -------------------------------------------------------------------------
    def setup(self):
        self.L.orc_setup(self.h)

    def run(self, nsteps, thermo_every=0):
        self.L.orc_run(self.h, nsteps, thermo_every, self.step)
        self.step += nsteps

    def compute_forces(self, eflag=1, vflag=1):
        self.L.orc_compute_forces(self.h, eflag, vflag)

    def thermo(self):
        out = np.zeros(10)
        self.L.orc_thermo(self.h, _p(out))
        return dict(pe=out[0], ke=out[1], temp=out[2], press=out[3], virial=out[4:10].copy())

    @property
    def nlocal(self):
        return self.L.orc_nlocal(self.h)

    @property
    def nghost(self):
        return self.L.orc_nghost(self.h)

    @property
    def ncalls(self):
        return self.L.orc_ncalls(self.h)

    @property
    def ndanger(self):
        return self.L.orc_ndanger(self.h)

    def bin_info(self):
        out = np.zeros(10, dtype=np.int32)
        self.L.orc_bin_info(self.h, _p(out))
        keys = ["nbinx", "nbiny", "nbinz", "mbinx", "mbiny", "mbinz", "mbinxlo", "mbinylo", "mbinzlo", "nstencil"]
        return dict(zip(keys, (int(v) for v in out)))

    def atoms(self):
        n = self.nlocal
        x = np.empty((n, 3)); v = np.empty((n, 3)); f = np.empty((n, 3))
        t = np.empty(n, dtype=np.int32); tag = np.empty(n, dtype=np.int32); img = np.empty(n, dtype=np.int32)
        self.L.orc_get_atoms(self.h, _p(x), _p(v), _p(f), _p(t), _p(tag), _p(img))
        return dict(x=x, v=v, f=f, type=t, tag=tag, image=img)

    def ghosts(self):
        n = self.nghost
        x = np.empty((n, 3)); tag = np.empty(n, dtype=np.int32)
        self.L.orc_get_ghosts(self.h, _p(x), _p(tag))
        return x, tag

    def pairs(self):
        """Stored half-list pairs as an (npairs, 3) int32 array of (tag_i, tag_j, code)."""
        n = self.L.orc_npairs(self.h)
        out = np.empty((n, 3), dtype=np.int32)
        self.L.orc_get_pairs(self.h, _p(out))
        return out


# ---------------------------------------------------------------------------------------------
# ready-made benchmark systems (bench/in.lj, bench/in.eam of the reference)
# ---------------------------------------------------------------------------------------------
def make_lj_system(nx=20, ny=None, nz=None, with_velocity=True):
    """bench/in.lj:3-27 : fcc 0.8442, T=1.44 seed 87287 loop geom, lj/cut 2.5, skin 0.3, every 20 check no."""
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    scale = lattice_scale_lj(0.8442)
    x, tag = create_fcc(nx, ny, nz, scale)
    type_ = np.ones(len(x), dtype=np.int32)
    mass = np.array([0.0, 1.0])
    v = velocity_create_geom(x, type_, mass, 1.44, 87287, "lj") if with_velocity else np.zeros_like(x)
    s = OracleSystem(1, "lj")
    s.set_box([0, 0, 0], [nx * scale, ny * scale, nz * scale])
    s.set_mass(1, 1.0)
    s.set_neighbor(0.3, every=20, delay=0, check=0)
    ones = np.ones((2, 2))
    s.pair_lj_cut(ones * 1.0, ones * 1.0, ones * 2.5, offset_flag=False)
    s.set_atoms(x, v, type_, tag)
    return s


def make_eam_system(potfile, nx=20, ny=None, nz=None, with_velocity=True):
    """bench/in.eam:3-31 : fcc 3.615, eam Cu_u3.eam, T=1600 seed 376847 loop geom, skin 1.0, every 1 delay 5 check yes."""
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    a = 3.615
    x, tag = create_fcc(nx, ny, nz, a)
    type_ = np.ones(len(x), dtype=np.int32)
    s = OracleSystem(1, "metal")
    s.set_dt(0.005)
    s.set_box([0, 0, 0], [nx * a, ny * a, nz * a])
    s.set_neighbor(1.0, every=1, delay=5, check=1)
    s.pair_eam({1: potfile})
    mass = np.array([0.0, s.mass(1)])
    v = velocity_create_geom(x, type_, mass, 1600.0, 376847, "metal") if with_velocity else np.zeros_like(x)
    s.set_atoms(x, v, type_, tag)
    return s


# ---------------------------------------------------------------------------------------------
# the unmodified reference (oracle/_ref)
# ---------------------------------------------------------------------------------------------
def ref_available() -> bool:
    return LMP_REF.exists() and os.access(LMP_REF, os.X_OK)


def run_ref_extract(out_prefix, pair_style, lammps_args, cwd=None, threads=1):
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    subprocess.check_call([str(REF_EXTRACT), str(out_prefix), pair_style] + list(lammps_args), cwd=cwd, env=env,
                          stdout=subprocess.DEVNULL)
    return read_ref_extract(out_prefix)


def read_ref_extract(out_prefix):
    meta = {}
    for line in Path(str(out_prefix) + ".meta").read_text().splitlines():
        k, *vals = line.split()
        meta[k] = [float(v) for v in vals] if len(vals) > 1 else float(vals[0])
    n = int(meta["nlocal"])
    raw = np.fromfile(str(out_prefix) + ".atoms.bin", dtype=np.uint8)
    off = 0
    tag = raw[off:off + 4 * n].view(np.int32); off += 4 * n
    typ = raw[off:off + 4 * n].view(np.int32); off += 4 * n
    x = raw[off:off + 24 * n].view(np.float64).reshape(n, 3); off += 24 * n
    v = raw[off:off + 24 * n].view(np.float64).reshape(n, 3); off += 24 * n
    f = raw[off:off + 24 * n].view(np.float64).reshape(n, 3)
    out = dict(meta=meta, tag=tag.copy(), type=typ.copy(), x=x.copy(), v=v.copy(), f=f.copy())
    npath = Path(str(out_prefix) + ".neigh.bin")
    if npath.exists():
        out["pairs"] = np.fromfile(npath, dtype=np.int32).reshape(-1, 3)
    return out
