"""MPM_Simulator (reference: mpm_solver/solver.py:9-177) on the B200 kernels.

Drop-in for `from mpm_solver.solver import *` in the reference's main.py: same constructor, same
`set_boundary_conditions / add_surface_collider / p2g2p / postprocess`, same `mpm_state` /
`mpm_model` attribute names, plus a `ti` shim because main.py obtains `ti` through the star import
and calls `ti.init(arch=ti.cuda)` (main.py:7,28).
"""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import _native
from .model import *            # noqa: F401,F403  (reference does the same star imports)
from .model import _Backend
from .utils import *            # noqa: F401,F403
from .utils import _stream, material_types
from .collider import *         # noqa: F401,F403
from .boundary_conditions import *  # noqa: F401,F403
from .fitting import FittingMixin, FittingModel, MPM_state_opt, _FittingBackend   # noqa: F401  (MPM_state_opt: reference name, model.py:136)


class _TiShim:
    """Stands in for the `taichi` module name main.py expects.  Only CUDA exists here."""
    cuda = "cuda"
    gpu = "cuda"
    cpu = "cpu"
    f32 = torch.float32
    i32 = torch.int32

    @staticmethod
    def init(arch="cuda", **kwargs):
        if arch != "cuda":
            raise RuntimeError("this mpm_solver has no CPU path: ti.init(arch=ti.cuda) only")
        if not torch.cuda.is_available():
            raise RuntimeError("no CUDA device visible")
        _native.lib()  # fail loudly now if the extension was not built


ti = _TiShim()


class MPM_Simulator(FittingMixin):
    def __init__(self, xyzs, covs, volumes, args, init_v=None, strict_q1=True, strict_q3=True, track_cov=False,
                 capacity=None, ids=None, n_global=None, strict_grad=True):
        self.n_particles = xyzs.shape[0]
        device = xyzs.device if xyzs.is_cuda else torch.device("cuda", torch.cuda.current_device())
        self.fitting = bool(getattr(args, "fitting", False))
        if self.fitting:
            # solver.py:14-15: the differentiable state (31-entry history, adjoint fields) and kernels (fitting.py, csrc/opt.cu);
            # the library context only provides the grid geometry and the stream plumbing here
            self.strict_grad = bool(strict_grad)
            self._backend = _FittingBackend(args.n_grid, args.grid_extent, device)
            self.mpm_model = FittingModel(self.n_particles, args, device, self._backend.ctx)
            self.mpm_state = MPM_state_opt(self.n_particles, xyzs, covs, volumes, args, init_v, device=device)
            self.time = 0.0
            self.collider_params = []
            self.particle_preprocess = []
            self.grid_postprocess = []
            self.init_particles = []
            return
        self._backend = _Backend(self.n_particles, args.n_grid, args.grid_extent, device, capacity=capacity, ids=ids,
                                 n_global=n_global)
        self.mpm_model = MPM_model(self.n_particles, args, _backend=self._backend)
        self.mpm_state = MPM_state(self.n_particles, xyzs, covs, volumes, args, _backend=self._backend)

        self.time = 0.0

        self.collider_params = []

        self.particle_preprocess = []
        self.grid_postprocess = []
        self.init_particles = []

        self._strict_q1 = strict_q1
        self._strict_q3 = strict_q3
        self._track_cov = track_cov
        self._sort_full = False
        self._sort_merge = False
        self._no_lag = False
        self._sync_flags()

    def load_particles(self, xyzs, covs, volumes):
        """Re-initialise the simulator in place from new particle data (same N): the constructor's state
        initialisation (model.py:101-115) without reallocating.  Registered boundary conditions stay."""
        if xyzs.shape[0] != self.n_particles:
            raise ValueError("load_particles: particle count must not change")
        self.mpm_state._init_particles(xyzs, covs, volumes)
        self.mpm_model._reinit_packed()
        for pp in self.init_particles:
            pp.apply(self.mpm_state, self.mpm_model)
            if pp.isMaterial == False:
                self.mpm_model.recompute_mu_lam()
                pp.applymu(self.mpm_state, self.mpm_model)
        self.time = 0.0
        self._sync_flags()

    # -- flags / parameters ------------------------------------------------------------------
    def _has_plastic(self):
        mat = self._backend.cur["meta"][: self.n_particles, 1]
        return bool(int(self._backend.read_small((mat != 0).any().reshape(1))[0])) if self.n_particles else False

    def _sync_flags(self):
        f = 0
        if self._strict_q1:
            f |= _native.STRICT_Q1
        if self._strict_q3:
            f |= _native.STRICT_Q3
        if self._track_cov:
            f |= _native.TRACK_COV
        if self._sort_full:
            f |= _native.SORT_FULL
        if self._sort_merge:
            f |= _native.SORT_MERGE
        if self._no_lag:
            f |= _native.NO_LAG
        if self._has_plastic():
            f |= _native.WRITE_FTRIAL  # F_trial != F after a return mapping; export needs F_trial
        self._backend.params.flags = f
        self.mpm_model._push()

    def set_parameters(self, params=None, **kw):
        """Update solver parameters from a dict (not in the reference; north_star asks for it).
        Keys: gravity, E, nu, yield_stress, hardening, xi, plastic_viscosity, softening, friction_angle,
        strict_q1, strict_q3, track_cov, sort_full (always the full radix sort, never the merge of the movers), sort_merge (the merge whenever valid,
        also for small clouds), no_lag (step() keeps the sort on the critical path of every substep instead of running it beside p2g
        on the previous substep's work list).  The physics keys take effect from the next substep on, as they would in the reference, whose
        p2g2p re-evaluates the stress from F_trial with the current parameters (solver.py:36)."""
        p = dict(params or {})
        p.update(kw)
        m = self.mpm_model
        for k, v in p.items():
            if k == "gravity":
                m.gravity = tuple(float(g) for g in v)
            elif k in ("hardening", "xi", "plastic_viscosity", "softening"):
                setattr(m, k, v)
            elif k == "friction_angle":
                m.friction_angle = float(v)
                sin_phi = math.sin(m.friction_angle / 180.0 * 3.141592653589793)
                m.alpha = math.sqrt(2.0 / 3.0) * 2.0 * sin_phi / (3.0 - sin_phi)
            elif k == "yield_stress":
                m.yield_stress.fill(float(v))
            elif k == "E":
                m._logE.fill_(math.log10(v))
            elif k == "nu":
                m._y.fill_(-math.log(0.49 / v - 1))
            elif k in ("strict_q1", "strict_q3", "track_cov", "sort_full", "sort_merge", "no_lag"):
                setattr(self, "_" + k, bool(v))
            else:
                raise KeyError(k)
        if "E" in p or "nu" in p:
            m.recompute_mu_lam()
        if any(k in p for k in ("hardening", "xi", "plastic_viscosity", "softening", "friction_angle", "strict_q1", "strict_q3")):
            self._backend.mark_stress_dirty()
        self._sync_flags()

    set_parameters_dict = set_parameters

    # -- the substep -------------------------------------------------------------------------
    def _ops_at(self, time) -> _native.SubstepOps:
        """active impulses and grid operations at `time`, in registration order (solver.py:31-46)"""
        ops = _native.SubstepOps()
        ni = 0
        for pp in self.particle_preprocess:
            if pp.isActive(time):
                if ni >= _native.MAX_IMPULSES:
                    raise _native.MpmkError("too many simultaneous impulses")
                ops.impulses[ni] = pp.as_impulse()
                ni += 1
        ops.n_impulses = ni
        no = 0
        for gp in self.grid_postprocess:
            if gp.isCollide or gp.isActive(time):
                if no >= _native.MAX_GRID_OPS:
                    raise _native.MpmkError("too many simultaneous grid operations")
                ops.ops[no] = gp.as_grid_op()
                no += 1
        ops.n_ops = no
        return ops

    def p2g2p(self, dt, s=None):
        if self.fitting:   # extra.py:200 calls p2g2p(dt, s) on the fitting simulator
            return self.p2g2p_forward(dt, 0 if s is None else s)
        b = self._backend
        ops = self._ops_at(self.time)
        b.flush_stress(dt)
        b.ctx.call("mpmk_substep", float(dt), C.byref(ops), _stream())
        b._ftrial_valid = bool(b.params.flags & _native.WRITE_FTRIAL)
        b.invalidate()
        self.time += dt

    def step(self, n_substeps, dt=None):
        """`n_substeps` calls of p2g2p(dt); stretches over which the active BC set is constant are
        replayed from a CUDA graph (mpmk_substeps).  `time` advances by the same double accumulation
        as the reference's loop (solver.py:52), so activity windows switch on the same substeps."""
        b = self._backend
        dt = float(self.mpm_model.args.substep_dt if dt is None else dt)
        if n_substeps > 0:
            b.flush_stress(dt)
        i = 0
        while i < n_substeps:
            ops = self._ops_at(self.time)
            raw = bytes(ops)
            t = self.time
            j = i
            while j < n_substeps and bytes(self._ops_at(t)) == raw:
                t += dt
                j += 1
            b.ctx.call("mpmk_substeps", dt, C.byref(ops), j - i, _stream())
            self.time = t
            i = j
        b._ftrial_valid = bool(b.params.flags & _native.WRITE_FTRIAL)
        b.invalidate()

    # -- registries: solver.py:110-129, :140-166 ---------------------------------------------
    def set_boundary_conditions(self, bc_args_arr, sim_args):
        for bc_args in bc_args_arr:
            bc_type = bc_args["type"]

            bc = boundaryConditionTypeCallBacks[bc_type](self.n_particles, bc_args, sim_args)

            if bc.type in preprocess_bc:
                self.particle_preprocess.append(bc)

            if bc.type in postprocess_bc:
                self.grid_postprocess.append(bc)

            if bc.type in init_bc:
                self.init_particles.append(bc)

        for pp in self.init_particles:
            pp.apply(self.mpm_state, self.mpm_model)
            if pp.isMaterial == False:
                self.mpm_model.recompute_mu_lam()
                pp.applymu(self.mpm_state, self.mpm_model)
        self._sync_flags()

    def add_surface_collider(
        self,
        point,
        normal,
        surface="sticky",  # accepted and ignored, as in the reference (solver.py:144)
        friction=0.0,
        start_time=0.0,    # ignored (solver.py:146)
        end_time=999.0,    # ignored (solver.py:147)
    ):
        point = list(point)
        normal_scale = 1.0 / math.sqrt(float(sum(x ** 2 for x in normal)))
        normal = list(normal_scale * x for x in normal)
        collider_param = MPM_Collider(point, normal, friction)
        self.collider_params.append(collider_param)
        cl = collideTypeCallBacks["ground"](collider_param.point, collider_param.normal, collider_param.friction)
        self.grid_postprocess.append(cl)

    def postprocess(self):
        """solver.py:135-137: compute_cov_from_F + compute_R_from_F, fused with the scatter to
        original particle order into persistent torch buffers (csrc/export.cu)."""
        b = self._backend
        b.export(want_R=True)
        b.cov_from_postprocess = True

    def export(self):
        """(xyz [N,3], cov [N,6], R [N,3,3]) CUDA tensors in original particle order, no host copy."""
        b = self._backend
        if not b.export_fresh:
            self.postprocess()
        return b.out_xyz, b.out_cov, b.out_R

    def export_world(self, scaling_modifier, pos_center, rotation_matrices=(), mode="render", render_scaling_modifier=1.0,
                     out=None):
        """Positions [N,3] and covariances [N,6] of the deformed Gaussians, original particle order, in ONE kernel
        (csrc/export.cu k_export_world) instead of postprocess() + Field.to_torch() + grid2world + the undo chain:
        mode="grid2world" gives what main.py:310-313 assigns to sim_means3D / sim_covs, mode="render" what
        render_frame (main.py:139-146) then hands to the rasteriser as means3D / cov3D_precomp."""
        from utils import transform_utils as T
        b = self._backend
        alpha, beta, gamma = T.world_affine(mode, scaling_modifier, pos_center, self.mpm_model.grid_extent, render_scaling_modifier)
        M = T.inverse_rotation_composite(rotation_matrices) if mode == "render" else [[1, 0, 0], [0, 1, 0], [0, 0, 1]]
        xf = _native.WorldXform()
        xf.alpha, xf.gamma = alpha, gamma
        for k in range(3):
            xf.beta[k] = float(beta[k])
            for c in range(3):
                xf.M[3 * k + c] = float(M[k][c])
        n = b.out_xyz.shape[0]
        if out is None:
            if getattr(b, "world_xyz", None) is None:
                b.world_xyz = torch.empty((n, 3), dtype=torch.float32, device=b.device)
                b.world_cov = torch.empty((n, 6), dtype=torch.float32, device=b.device)
            out = (b.world_xyz, b.world_cov)
        b.ctx.call("mpmk_export_world", b.init_cov.data_ptr(), C.byref(xf), out[0].data_ptr(), out[1].data_ptr(), _stream())
        return out

    def counters(self):
        c = _native.Counters()
        self._backend.ctx.call("mpmk_read_counters", C.byref(c), _stream())
        return {k: getattr(c, k) for k, _ in c._fields_}

    # differentiable API of the reference (solver.py:54-108,131-133,168-177): FittingMixin (fitting.py), needs args.fitting
