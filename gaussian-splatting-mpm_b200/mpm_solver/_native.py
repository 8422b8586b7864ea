"""ctypes binding of lib/libmpmk.so (ABI: include/mpmk.h).

There is no CPU fallback: if the CUDA library is missing or the device is not a
CUDA device every entry point raises.
"""
from __future__ import annotations

import ctypes as C
import os

_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# MPMK_LIB selects another build of the same library (kernel tuning experiments); default in-tree build
LIB_PATH = os.environ.get("MPMK_LIB") or os.path.join(_ROOT, "lib", "libmpmk.so")

MAT_JELLY, MAT_METAL, MAT_SAND, MAT_FOAM, MAT_FLUID, MAT_PLASTICINE = range(6)
STRICT_Q1, STRICT_Q3, WRITE_FTRIAL, TRACK_COV, SORT_FULL, SORT_MERGE, NO_LAG = 1, 2, 4, 8, 16, 64, 128
OP_FIXED_CUBE, OP_COLLIDER = 0, 1
MAX_GRID_OPS, MAX_IMPULSES = 16, 8

SET_FIELDS = ("xm", "vv", "c0", "c1", "cs", "s1", "f0", "f1", "f2", "meta", "cv0", "cv1")


class ParticleSet(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in SET_FIELDS]


class Bindings(C.Structure):
    _fields_ = [("set", ParticleSet * 2), ("ft0", C.c_void_p), ("ft1", C.c_void_p), ("ft2", C.c_void_p),
                ("keys", C.c_void_p), ("grid_in", C.c_void_p), ("grid_out", C.c_void_p)]


class Params(C.Structure):
    _fields_ = [("gravity", C.c_float * 3), ("alpha", C.c_float), ("plastic_viscosity", C.c_float),
                ("xi", C.c_float), ("softening", C.c_float), ("hardening", C.c_int32), ("flags", C.c_uint32)]


class GridOp(C.Structure):
    _fields_ = [("type", C.c_int32), ("a", C.c_float * 3), ("b", C.c_float * 3), ("c", C.c_float)]


class Impulse(C.Structure):
    _fields_ = [("center", C.c_float * 3), ("size", C.c_float * 3), ("force", C.c_float * 3),
                ("substep_dt", C.c_float)]


class SubstepOps(C.Structure):
    _fields_ = [("n_ops", C.c_int32), ("ops", GridOp * MAX_GRID_OPS), ("n_impulses", C.c_int32),
                ("impulses", Impulse * MAX_IMPULSES)]


class WorldXform(C.Structure):
    _fields_ = [("alpha", C.c_float), ("gamma", C.c_float), ("beta", C.c_float * 3), ("M", C.c_float * 9)]


class Counters(C.Structure):
    _fields_ = [("n_items", C.c_uint32), ("n_active_blocks", C.c_uint32), ("n_tiles", C.c_uint32),
                ("n_clamped", C.c_uint32), ("n_crossers", C.c_uint32)]


class MpmkError(RuntimeError):
    pass


NCCL_ID_BYTES = 128


def nccl_library_path():
    """Path of the libnccl.so.2 this process's torch uses (the data path dlopens the same library)."""
    try:
        with open("/proc/self/maps") as f:
            for line in f:
                if "libnccl.so" in line:
                    return line.split()[-1]
    except OSError:
        pass
    import glob
    import sysconfig
    hits = glob.glob(os.path.join(sysconfig.get_paths()["purelib"], "nvidia", "nccl", "lib", "libnccl.so*"))
    return hits[0] if hits else "libnccl.so.2"


_lib = None


def lib():
    """Load libmpmk.so; raises if it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MpmkError(f"{LIB_PATH} is missing: build it with gaussian-splatting-mpm_b200/build.sh "
                        "(python -c 'import __graft_entry__ as g; g.build()'); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, i, f, i64 = C.c_void_p, C.c_int, C.c_float, C.c_int64
    sig = {
        "mpmk_version": (i, []),
        "mpmk_create": (i, [C.POINTER(vp), i, i, C.c_double, i64]),
        "mpmk_destroy": (i, [vp]),
        "mpmk_last_error": (C.c_char_p, [vp]),
        "mpmk_bind": (i, [vp, C.POINTER(Bindings)]),
        "mpmk_set_params": (i, [vp, C.POINTER(Params)]),
        "mpmk_set_count": (i, [vp, i64, i]),
        "mpmk_current_set": (i, [vp]),
        "mpmk_count": (i64, [vp]),
        "mpmk_compute_mu_lam": (i, [vp, i64, vp, vp, vp, vp, vp]),
        "mpmk_particle_volume": (i, [vp, i64, vp, f, vp, vp, vp]),
        "mpmk_particle_volume_noalloc": (i, [vp, i64, vp, f, vp, vp]),
        "mpmk_compute_keys": (i, [vp, vp]),
        "mpmk_constitutive": (i, [vp, f, vp]),
        "mpmk_constitutive_ex": (i, [vp, f, i, i, vp]),
        "mpmk_sort": (i, [vp, vp]),
        "mpmk_get_sort_result": (i, [vp, vp, vp, vp]),
        "mpmk_get_last_sort": (i, [vp, vp, vp, vp, vp]),
        "mpmk_debug_trace": (i, [vp, vp]),
        "mpmk_sort_info": (i, [vp, C.POINTER(i), C.POINTER(i64), vp]),
        "mpmk_prepare_grid": (i, [vp, vp]),
        "mpmk_p2g": (i, [vp, f, C.POINTER(SubstepOps), vp]),
        "mpmk_grid_update": (i, [vp, f, C.POINTER(SubstepOps), vp]),
        "mpmk_g2p": (i, [vp, f, vp]),
        "mpmk_substep": (i, [vp, f, C.POINTER(SubstepOps), vp]),
        "mpmk_substeps": (i, [vp, f, C.POINTER(SubstepOps), i, vp]),
        "mpmk_export": (i, [vp, vp, vp, vp, vp, vp]),
        "mpmk_export_world": (i, [vp, vp, C.POINTER(WorldXform), vp, vp, vp]),
        "mpmk_read_counters": (i, [vp, C.POINTER(Counters), vp]),
        "mpmk_opt_stress": (i, [vp, i64, vp, vp, vp, vp, vp]),
        "mpmk_opt_stress_grad": (i, [vp, i64, vp, vp, vp, vp, vp, vp, vp, vp]),
        "mpmk_opt_p2g": (i, [vp, i64, vp, vp, vp, vp, vp, vp, vp, vp, f, vp]),
        "mpmk_opt_p2g_grad": (i, [vp, i64, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, f, vp]),
        "mpmk_opt_grid": (i, [vp, vp, vp, vp, f, vp, vp, vp, vp]),
        "mpmk_opt_grid_grad": (i, [vp, vp, vp, vp, vp, vp, vp]),
        "mpmk_opt_g2p": (i, [vp, i64, vp, vp, vp, vp, vp, vp, vp, f, vp]),
        "mpmk_opt_g2p_grad": (i, [vp, i64, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, f, vp]),
        "mpmk_opt_cov": (i, [vp, i64, vp, vp, vp, vp]),
        "mpmk_opt_cov_grad": (i, [vp, i64, vp, vp, vp, vp, vp]),
        "mpmk_opt_mu_lam_grad": (i, [vp, i64, vp, vp, vp, vp, vp, vp, vp]),
        "mpmk_opt_learn": (i, [vp, i64, vp, vp, vp, vp, vp]),
        "mpmk_launch_count": (i64, [vp]),
        "mpmk_mg_init": (i, [vp, C.c_char_p, vp, i, i, i]),
        "mpmk_mg_finalize": (i, [vp]),
        "mpmk_mg_set_slab": (i, [vp, i, i, i, i, i, vp, vp]),
        "mpmk_mg_set_window": (i, [vp, i, i]),
        "mpmk_mg_peer_error": (i, [vp, C.POINTER(C.c_uint32), vp]),
        "mpmk_mg_halo_exchange": (i, [vp, vp]),
        "mpmk_mg_peer_alloc": (i, [vp, i64, vp, i]),
        "mpmk_mg_peer_alloc_ex": (i, [vp, i64, i64, vp, i]),
        "mpmk_mg_ll_send": (i, [vp, i, vp, i64, C.c_uint32, vp]),
        "mpmk_mg_ll_recv": (i, [vp, i, vp, i64, C.c_uint32, vp]),
        "mpmk_mg_peer_open": (i, [vp, vp, vp]),
        "mpmk_mg_peer_enabled": (i, [vp]),
        "mpmk_mg_peer_close": (i, [vp]),
        "mpmk_mg_set_count": (i, [vp, i64]),
        "mpmk_read_small": (i, [vp, vp, vp, i, vp]),
        "mpmk_mg_scan": (i, [vp, vp, vp, vp, vp]),
        "mpmk_mg_pack": (i, [vp, vp, i64, vp, vp]),
        "mpmk_mg_unpack": (i, [vp, vp, i64, i64, vp, vp]),
        "mpmk_mg_close_holes": (i, [vp, vp, i64, i64, vp, i64, vp]),
    }
    L.mpmk_mg_unique_id.restype = i
    L.mpmk_mg_unique_id.argtypes = [C.c_char_p, vp, i]
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


class Context:
    """Owns one mpmk_ctx; methods raise MpmkError with the library's message on failure."""

    def __init__(self, device_index: int, n_grid: int, grid_extent: float, capacity: int):
        self.L = lib()
        self.h = C.c_void_p()
        rc = self.L.mpmk_create(C.byref(self.h), int(device_index), int(n_grid), float(grid_extent), int(capacity))
        if rc != 0:
            msg = self.L.mpmk_last_error(self.h).decode() if self.h else "mpmk_create failed"
            if self.h:
                self.L.mpmk_destroy(self.h)
                self.h = C.c_void_p()
            raise MpmkError(f"mpmk_create: {msg} (code {rc})")

    def call(self, name, *args):
        rc = getattr(self.L, name)(self.h, *args)
        if rc != 0:
            raise MpmkError(f"{name}: {self.L.mpmk_last_error(self.h).decode()} (code {rc})")

    def close(self):
        if getattr(self, "h", None):
            self.L.mpmk_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass
