"""ctypes/numpy front end of the CPU oracle (oracle/mpm_oracle.c).

TEST INFRASTRUCTURE ONLY -- see the header of mpm_oracle.c.  PARITY UNPINNED: the
reference has no tests or golden vectors for this path and Taichi is not
installable here; the oracle is a line-by-line restatement.

`OracleSim` mirrors the reference's `MPM_Simulator` (mpm_solver/solver.py:9-166) on
numpy arrays in the reference's own field layout (mpm_solver/model.py:77-127), so
parity tests read like calls into the reference.
"""
from __future__ import annotations

import ctypes as C
import math
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))

MATERIAL_TYPES = {  # mpm_solver/utils.py:5-10 (+ fluid/plasticine: SURVEY.md 9.5, unpinned)
    "jelly": 0, "metal": 1, "sand": 2, "foam": 3, "fluid": 4, "plasticine": 5,
}
STRICT_Q1 = 1
STRICT_Q3 = 2


class OrcParams(C.Structure):
    _fields_ = [("alpha", C.c_float), ("plastic_viscosity", C.c_float), ("xi", C.c_float),
                ("softening", C.c_float), ("hardening", C.c_int), ("flags", C.c_int)]


def build(force: bool = False) -> None:
    """Compile both oracle libraries in place (gcc; a few seconds)."""
    need = force or not all(os.path.exists(os.path.join(_HERE, f))
                            for f in ("libmpm_oracle.so", "libmpm_oracle_omp.so"))
    if need:
        subprocess.check_call(["make", "-C", _HERE, "-B", "-s"])


_fp = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
_libs = {}


def lib(omp: bool = False):
    if omp in _libs:
        return _libs[omp]
    build()
    L = C.CDLL(os.path.join(_HERE, "libmpm_oracle_omp.so" if omp else "libmpm_oracle.so"))
    i, f, vp = C.c_int, C.c_float, C.c_void_p
    PP = C.POINTER(OrcParams)
    L.orc_svd3.argtypes = [_fp, _fp, _fp, _fp]
    L.orc_compute_stress_from_F_trial.argtypes = [i, _fp, _fp, _fp, _fp, _fp, _fp, _fp, vp, f, PP]
    L.orc_cell_keys.argtypes = [i, i, f, _fp, vp, _u32p]
    L.orc_stable_sort.argtypes = [i, _u32p, _u32p, _u32p]
    L.orc_p2g.argtypes = [i, i, f, f, f, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp]
    L.orc_grid_normalization_and_gravity.argtypes = [i, f, _fp, _fp, _fp, _fp]
    L.orc_fixed_cube.argtypes = [i, f, _fp, _fp, _fp]
    L.orc_collide.argtypes = [i, f, _fp, _fp, f, _fp]
    L.orc_impulse.argtypes = [i, _fp, _fp, _fp, _fp, _fp, _fp, f]
    L.orc_g2p.argtypes = [i, i, f, f, f, _fp, _fp, _fp, _fp, _fp, vp, _fp]
    L.orc_compute_cov_from_F.argtypes = [i, _fp, _fp, _fp]
    L.orc_compute_R_from_F.argtypes = [i, _fp, _fp]
    L.orc_compute_mu_lam_from_E_nu.argtypes = [i, _fp, _fp, _fp, _fp]
    L.orc_compute_mass_from_vol_density.argtypes = [i, _fp, _fp, _fp]
    L.orc_particle_volume.argtypes = [i, _fp, i, f, _fp]
    L.orc_reset_grid.argtypes = [i, _fp, _fp, _fp]
    L.orc_num_threads.restype = i
    L.orc_set_num_threads.argtypes = [i]
    L.orc_set_num_threads.restype = None
    for name in dir(L):
        pass
    _libs[omp] = L
    return L


def _f32(a):
    return np.ascontiguousarray(np.asarray(a, dtype=np.float32))


def svd3(A):
    """ti.svd restatement on one 3x3 (returns U, sigma[3], V)."""
    A = _f32(A).reshape(9)
    U = np.empty(9, np.float32); s = np.empty(3, np.float32); V = np.empty(9, np.float32)
    lib().orc_svd3(A, U, s, V)
    return U.reshape(3, 3), s, V.reshape(3, 3)


def cell_keys(xyz, n_grid, inv_dx):
    """(base[N,3] int32, key[N] uint32) of this build's sort key; base per utils.py:94-95."""
    xyz = _f32(xyz)
    n = xyz.shape[0]
    base = np.empty((n, 3), np.int32); keys = np.empty(n, np.uint32)
    lib().orc_cell_keys(n, int(n_grid), np.float32(inv_dx), xyz, base.ctypes.data, keys)
    return base, keys


def stable_sort(keys):
    keys = np.ascontiguousarray(keys, dtype=np.uint32)
    sk = np.empty_like(keys); perm = np.empty_like(keys)
    lib().orc_stable_sort(keys.shape[0], keys, sk, perm)
    return sk, perm


def particle_volume(pos, n_grid, grid_extent):
    """internel_filling/filling.py:27-42 get_particle_volume(uniform=False)."""
    pos = _f32(pos)
    vol = np.empty(pos.shape[0], np.float32)
    lib().orc_particle_volume(pos.shape[0], pos, int(n_grid), np.float32(grid_extent / n_grid), vol)
    return vol


def default_alpha() -> float:
    # mpm_solver/model.py:49-51 (Python-scope doubles)
    sin_phi = math.sin(25.0 / 180.0 * 3.141592653589793)
    return math.sqrt(2.0 / 3.0) * 2.0 * sin_phi / (3.0 - sin_phi)


class _BC:
    """boundary_conditions.py:7-31 BasicBC bookkeeping (host side)."""

    def __init__(self, bc_args, substep_dt):
        self.substep_dt = substep_dt
        self.id = bc_args["id"]
        self.type = bc_args["type"]
        self.start_time = bc_args["start_time"]
        self.end_time = bc_args["start_time"] + substep_dt * bc_args["num_dt"]
        self.center = _f32(bc_args["center"])
        self.size = _f32(bc_args["size"])
        self.force = _f32(bc_args["force"]) if "force" in bc_args else None
        self.isCollide = False

    def isActive(self, time):
        return time >= self.start_time and time < self.end_time


class _Collider:
    def __init__(self, point, normal, friction):
        self.point = _f32(point); self.normal = _f32(normal); self.friction = float(friction)
        self.isCollide = True


class OracleSim:
    """numpy mirror of MPM_Simulator / MPM_model / MPM_state.

    sort=True additionally applies this build's ordering contract: at the start of
    every substep the particles are stably sorted by cell key relative to their
    current physical order and physically reordered (`order` maps physical slot ->
    original particle id).  sort=False is the reference as written (no sort).
    """

    def __init__(self, xyzs, covs, volumes, args, omp=False, strict_q1=True, strict_q3=True,
                 sort=True, track_cov=True):
        self.L = lib(omp)
        self.omp = omp
        self.sort = sort
        self.track_cov = track_cov
        n = self.n_particles = int(np.asarray(xyzs).shape[0])
        self.n_grid = int(args.n_grid)
        self.grid_extent = float(args.grid_extent)
        self.dx = np.float32(args.grid_extent / args.n_grid)          # model.py:15
        self.inv_dx = np.float32(args.n_grid / args.grid_extent)      # model.py:16
        m = MATERIAL_TYPES.get(args.material, -1)
        if m < 0:
            raise TypeError("Material not supported yet")              # model.py:30
        self.gravity = _f32(args.gravity)
        self.material = np.full(n, m, np.float32)                      # model.py:25 (f32 field)
        self.logE = np.full(n, math.log10(args.E), np.float32)
        self.y = np.full(n, -math.log(0.49 / args.nu - 1), np.float32)
        self.mu = np.empty(n, np.float32); self.lam = np.empty(n, np.float32)
        self.L.orc_compute_mu_lam_from_E_nu(n, self.logE, self.y, self.mu, self.lam)
        self.yield_stress = np.full(n, 0.005, np.float32)              # model.py:54-55
        self.params = OrcParams(alpha=default_alpha(), plastic_viscosity=0.008, xi=1.0, softening=1.0,
                                hardening=1,
                                flags=(STRICT_Q1 if strict_q1 else 0) | (STRICT_Q3 if strict_q3 else 0))
        # state, model.py:82-121
        self.particle_vol = _f32(volumes).reshape(n).copy()
        self.particle_density = np.full(n, args.density, np.float32)
        self.particle_mass = np.empty(n, np.float32)
        self.L.orc_compute_mass_from_vol_density(n, self.particle_density, self.particle_vol, self.particle_mass)
        self.particle_xyz = _f32(xyzs).reshape(n, 3).copy()
        self.particle_vel = np.zeros((n, 3), np.float32)
        eye = np.tile(np.eye(3, dtype=np.float32), (n, 1, 1))
        self.particle_F = eye.copy()
        self.particle_F_trial = eye.copy()
        self.particle_stress = np.zeros((n, 3, 3), np.float32)
        self.particle_C = np.zeros((n, 3, 3), np.float32)
        self.particle_pressure = np.zeros(n, np.float32)
        self.particle_init_cov = _f32(covs).reshape(n * 6).copy()
        self.particle_cov = self.particle_init_cov.copy()
        self.particle_R = np.zeros((n, 3, 3), np.float32)
        g = self.n_grid
        self.grid_mass = np.zeros((g, g, g), np.float32)
        self.grid_v_in = np.zeros((g, g, g, 3), np.float32)
        self.grid_v_out = np.zeros((g, g, g, 3), np.float32)
        self.order = np.arange(n, dtype=np.uint32)
        self.last_keys = None
        self.last_perm = None
        self.time = 0.0
        self.particle_preprocess = []
        self.grid_postprocess = []
        self.substep_dt = float(getattr(args, "substep_dt", 0.0))

    # -- registries: solver.py:110-129, :140-166 ---------------------------------------------
    def set_boundary_conditions(self, bc_args_arr, sim_args):
        for a in bc_args_arr:
            t = a["type"]
            if t not in ("fixed_cube", "impulse", "modify_material", "additional_params"):
                raise KeyError(t)
            bc = _BC(a, sim_args.substep_dt)
            if t == "impulse":
                self.particle_preprocess.append(bc)
            elif t == "fixed_cube":
                self.grid_postprocess.append(bc)
            else:
                inside = np.all(np.abs(self.particle_xyz - bc.center) < bc.size, axis=1)
                if t == "modify_material":      # boundary_conditions.py:81-85
                    self.material[inside] = np.float32(a["material"])
                else:                           # additional_params, intended behaviour (Q6, unpinned)
                    self.logE[inside] = np.float32(math.log10(a["E"]))
                    self.y[inside] = np.float32(-math.log(0.49 / a["nu"] - 1))
                    self.particle_density[inside] = np.float32(a["density"])
                    self.L.orc_compute_mu_lam_from_E_nu(self.n_particles, self.logE, self.y, self.mu, self.lam)
                    if a["mu"] != 1000:
                        self.mu[inside] = np.float32(a["mu"])
                    # boundary_conditions.py:62 only assigns particle_density; particle_mass was computed in
                    # MPM_state.__init__ (model.py:107) and is NOT recomputed -- the override never reaches the mass

    def add_surface_collider(self, point, normal, surface="sticky", friction=0.0, start_time=0.0, end_time=999.0):
        s = 1.0 / math.sqrt(float(sum(x ** 2 for x in normal)))
        self.grid_postprocess.append(_Collider(list(point), [s * x for x in normal], friction))

    # -- ordering contract of this build -----------------------------------------------------
    _PER_PARTICLE = ("material", "logE", "y", "mu", "lam", "yield_stress", "particle_vol", "particle_density",
                     "particle_mass", "particle_xyz", "particle_vel", "particle_F", "particle_F_trial",
                     "particle_stress", "particle_C", "particle_pressure", "particle_R")

    def _sort_and_reorder(self):
        _, keys = cell_keys(self.particle_xyz, self.n_grid, self.inv_dx)
        sk, perm = stable_sort(keys)
        self.last_keys, self.last_perm = sk, perm
        for name in self._PER_PARTICLE:
            setattr(self, name, np.ascontiguousarray(getattr(self, name)[perm]))
        n = self.n_particles
        self.particle_init_cov = np.ascontiguousarray(self.particle_init_cov.reshape(n, 6)[perm]).reshape(-1)
        self.particle_cov = np.ascontiguousarray(self.particle_cov.reshape(n, 6)[perm]).reshape(-1)
        self.order = np.ascontiguousarray(self.order[perm])

    # -- the substep: solver.py:27-52 --------------------------------------------------------
    def compute_stress(self, dt):
        self.L.orc_compute_stress_from_F_trial(
            self.n_particles, self.material, self.mu, self.lam, self.yield_stress,
            self.particle_F_trial.reshape(-1), self.particle_F.reshape(-1), self.particle_stress.reshape(-1),
            self.particle_pressure.ctypes.data, np.float32(dt), C.byref(self.params))

    def p2g(self, dt):
        self.L.orc_p2g(self.n_particles, self.n_grid, self.dx, self.inv_dx, np.float32(dt),
                       self.particle_xyz.reshape(-1), self.particle_vel.reshape(-1), self.particle_C.reshape(-1),
                       self.particle_stress.reshape(-1), self.particle_vol, self.particle_mass,
                       self.grid_v_in.reshape(-1), self.grid_mass.reshape(-1))

    def grid_ops(self, dt):
        self.L.orc_grid_normalization_and_gravity(self.n_grid, np.float32(dt), self.gravity,
                                                  self.grid_mass.reshape(-1), self.grid_v_in.reshape(-1),
                                                  self.grid_v_out.reshape(-1))
        for gp in self.grid_postprocess:
            if gp.isCollide:
                self.L.orc_collide(self.n_grid, self.dx, gp.point, gp.normal, np.float32(gp.friction),
                                   self.grid_v_out.reshape(-1))
            elif gp.isActive(self.time):
                self.L.orc_fixed_cube(self.n_grid, self.dx, gp.center, gp.size, self.grid_v_out.reshape(-1))

    def g2p(self, dt):
        self.L.orc_g2p(self.n_particles, self.n_grid, self.dx, self.inv_dx, np.float32(dt),
                       self.particle_xyz.reshape(-1), self.particle_vel.reshape(-1), self.particle_C.reshape(-1),
                       self.particle_F.reshape(-1), self.particle_F_trial.reshape(-1),
                       self.particle_cov.ctypes.data if self.track_cov else None, self.grid_v_out.reshape(-1))

    def p2g2p(self, dt):
        if self.sort:
            self._sort_and_reorder()
        self.L.orc_reset_grid(self.n_grid, self.grid_mass.reshape(-1), self.grid_v_in.reshape(-1),
                              self.grid_v_out.reshape(-1))
        for pp in self.particle_preprocess:
            if pp.isActive(self.time):
                self.L.orc_impulse(self.n_particles, self.particle_xyz.reshape(-1), self.particle_vel.reshape(-1),
                                   self.particle_mass, pp.center, pp.size, pp.force, np.float32(pp.substep_dt))
        self.compute_stress(dt)
        self.p2g(dt)
        self.grid_ops(dt)
        self.g2p(dt)
        self.time += dt

    def postprocess(self):  # solver.py:135-137
        self.L.orc_compute_cov_from_F(self.n_particles, self.particle_F_trial.reshape(-1),
                                      self.particle_init_cov, self.particle_cov)
        self.L.orc_compute_R_from_F(self.n_particles, self.particle_F_trial.reshape(-1),
                                    self.particle_R.reshape(-1))

    # -- original-order views (what FieldView.to_torch() returns in the product) --------------
    def in_original_order(self, name):
        a = getattr(self, name)
        n = self.n_particles
        flat6 = name in ("particle_cov", "particle_init_cov")
        if flat6:
            a = a.reshape(n, 6)
        out = np.empty_like(a)
        out[self.order] = a
        return out.reshape(-1) if flat6 else out
