"""ctypes binding of include/b200md.h.  Every entry point of the header is declared here; a missing
library or symbol raises immediately (the product never degrades to a CPU path)."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

from .build import LIB, build

OK = 0
ERR_NAMES = {1: "NONFINITE", 2: "NEIGH_OVERFLOW", 3: "LOST_ATOMS", 4: "CUDA", 5: "ARG", 6: "DOMAIN"}
NCCL_ID_BYTES = 128
IPC_HANDLE_BYTES = 64   # B200MD_IPC_HANDLE_BYTES

_vp, _i, _d, _i64 = C.c_void_p, C.c_int, C.c_double, C.c_int64

# name -> (restype, argtypes); mirrors include/b200md.h one to one
PROTOTYPES = {
    "b200md_nccl_unique_id": (_i, [_vp]),
    "b200md_create": (_i, [C.POINTER(_vp), _i, _i, _i, _vp]),
    "b200md_destroy": (_i, [_vp]),
    "b200md_last_error": (C.c_char_p, [_vp]),
    "b200md_memory_usage": (_d, [_vp]),
    "b200md_set_units": (_i, [_vp, _d, _d, _d, _d]),
    "b200md_set_box": (_i, [_vp, _vp, _vp, _vp, _vp]),
    "b200md_set_mass": (_i, [_vp, _i, _vp]),
    "b200md_upload_atoms": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp]),
    "b200md_nlocal": (_i, [_vp]),
    "b200md_nghost": (_i, [_vp]),
    "b200md_download_atoms": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "b200md_download_atoms_by_tag": (_i, [_vp, _i, _vp, _i, _vp, _vp, _vp, _vp]),
    "b200md_download_all": (_i, [_vp, _vp, _vp]),
    "b200md_halo_put_export": (_i, [_vp, C.c_int64, _vp]),
    "b200md_halo_put_import": (_i, [_vp, _i, _vp, _vp]),
    "b200md_halo_put_count": (C.c_int64, [_vp]),
    "b200md_lj_init": (_i, [_vp, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "b200md_eam_init": (_i, [_vp, _i, _i, _i, _d, _d, _d, _d, _i, _vp, _i, _vp, _i, _vp, _vp, _vp, _vp]),
    "b200md_eam_init_funcfl": (_i, [_vp, _i, _i, _vp, _vp, _vp, _vp]),
    "b200md_eam_init_setfl": (_i, [_vp, _i, C.c_char_p, _i, _vp, _vp, _vp]),
    "b200md_eam_download": (_i, [_vp, _vp, _vp]),
    "b200md_neigh_config": (_i, [_vp, _i, _vp, _d, _d, _i, _i, _i, _d]),
    "b200md_neigh_decide": (_i, [_vp, _vp]),
    "b200md_neigh_build": (_i, [_vp]),
    "b200md_neigh_export": (_i, [_vp, _vp, _vp, _vp]),
    "b200md_neigh_stats": (_i, [_vp, _vp, _vp, _vp]),
    "b200md_halo_forward": (_i, [_vp]),
    "b200md_plan_peers": (_i, [_vp, _vp, _i, _vp, _vp]),
    "b200md_pair_compute": (_i, [_vp, _i, _i, _vp, _vp]),
    "b200md_pair_peratom": (_i, [_vp]),
    "b200md_peratom_download": (_i, [_vp, _vp, _vp]),
    "b200md_nve_initial": (_i, [_vp, _d, _d]),
    "b200md_nve_final": (_i, [_vp, _d]),
    "b200md_thermo": (_i, [_vp, _d, _vp]),
    "b200md_last_ev": (_i, [_vp, _vp, _vp]),
    "b200md_setup": (_i, [_vp]),
    "b200md_run": (_i, [_vp, _i, _d, _i]),
    "b200md_timestep": (_i64, [_vp]),
    "b200md_set_timestep": (_i, [_vp, _i64]),
    "b200md_step_host": (_i, [_vp, _d, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "b200md_host_atoms": (_i, [_vp, _i, _i, _vp, _vp, _vp, _vp, _i]),
    "b200md_host_forces": (_i, [_vp, _vp, _i]),
    "b200md_host_eam_rho": (_i, [_vp, _i, _vp, _vp, _vp]),
    "b200md_host_eam_force": (_i, [_vp, _vp, _i, _i, _vp, _vp]),
    "b200md_set_bonds": (_i, [_vp, _i, _vp, _i, _vp, _vp, _vp, _i]),
    "b200md_bond_fene": (_i, [_vp, _i, _vp, _vp, _vp, _vp]),
    "b200md_thermo_bond": (_i, [_vp, _vp, _vp]),
    "b200md_fix_langevin": (_i, [_vp, _d, _d, _d, _i, _d]),
    "b200md_langevin_order": (_i, [_vp, _i, _vp, _i, _d]),
    "b200md_run_bounds": (_i, [_vp, _i64, _i64]),
    "b200md_ranmars": (_i, [_i, _i, _vp]),
    "b200md_langevin_rng": (_i, [_vp, _i]),
    "b200md_langevin_counter_uniforms": (_i, [_i, _i, _i64, _vp]),
    "b200md_philox4x32_10": (_i, [_vp, _vp, _vp]),
    "b200md_timers": (_i, [_vp, _vp, _vp, _i]),
    "b200md_profile": (_i, [_vp, _i]),
    "b200md_kernel_times": (_i, [_vp, _vp, _vp, _i]),
    "b200md_stream": (_vp, [_vp]),
    "b200md_tune": (_i, [_vp, C.c_char_p, _i]),
    "b200md_host_special": (_i, [_vp, _i, _i, _vp, _i, _vp, _vp, _vp]),
    "b200md_build_bins_per_block": (_i, [_vp]),
    "b200md_ke_noalloc": (_i, [_i, _i, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "b200md_ke_noalloc_max_atoms": (_i, []),
}

_lib = None


class B200MDError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"b200md error {code} ({ERR_NAMES.get(code, '?')}): {text}")
        self.code = code


def load(rebuild: bool = False) -> C.CDLL:
    """dlopen libb200md.so (building it first when stale) and declare every prototype."""
    global _lib
    if _lib is None or rebuild:
        path = build(force=rebuild)
        if not Path(path).exists():
            raise ImportError(f"{path} is missing: the CUDA library must be built (no CPU fallback)")
        L = C.CDLL(str(path))
        for name, (res, args) in PROTOTYPES.items():
            fn = getattr(L, name)  # AttributeError if the symbol is missing
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def ptr(a):
    """void* of a C-contiguous numpy array (or None), or of a torch tensor (host or device)."""
    if a is None:
        return None
    if isinstance(a, np.ndarray):
        assert a.flags["C_CONTIGUOUS"]
        return a.ctypes.data_as(C.c_void_p)
    if hasattr(a, "data_ptr"):
        return C.c_void_p(a.data_ptr())
    raise TypeError(type(a))


class Context:
    """Owner of one b200md_ctx (one GPU, one rank)."""

    def __init__(self, device: int = 0, nranks: int = 1, rank: int = 0, nccl_id: bytes | None = None):
        self.L = load()
        self.h = _vp()
        idbuf = C.create_string_buffer(nccl_id, NCCL_ID_BYTES) if nccl_id is not None else None
        rc = self.L.b200md_create(C.byref(self.h), device, nranks, rank, idbuf)
        if rc != OK:
            text = self.L.b200md_last_error(self.h if self.h else None).decode()
            if self.h:
                self.L.b200md_destroy(self.h)
                self.h = _vp()
            raise B200MDError(rc, text)

    def call(self, name, *args):
        rc = getattr(self.L, name)(self.h, *args)
        if rc != OK:
            raise B200MDError(rc, self.L.b200md_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.L.b200md_destroy(self.h)
            self.h = _vp()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def nccl_unique_id() -> bytes:
    L = load()
    buf = C.create_string_buffer(NCCL_ID_BYTES)
    rc = L.b200md_nccl_unique_id(buf)
    if rc != OK:
        raise B200MDError(rc, L.b200md_last_error(None).decode())
    return buf.raw
