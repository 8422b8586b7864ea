"""The differentiable ("fitting") path of the reference -- MPM_state_opt (mpm_solver/model.py:136-223) and the
p2g2p_forward / p2g2p_backward / learn / postprocess_forward / postprocess_backward / clear_grads / set_bc_ground_only
methods of MPM_Simulator (mpm_solver/solver.py:54-108,131-133,168-177), as extra.py:164-256 drives them -- on the
hand-written forward + adjoint kernels of csrc/opt.cu (C ABI: include/mpmk.h, mpmk_opt_*).  No Taichi, no autograd:
torch only owns the memory.

Field objects keep the reference's surface: `.to_torch()`, `.from_torch()`, `.fill()`, and `.grad` (the adjoint
field Taichi allocates with needs_grad=True).  History-indexed fields have the reference's shapes:
particle_xyz / particle_vel (31, N, 3), particle_F / particle_C / particle_stress (31, N, 3, 3).
"""
from __future__ import annotations

import torch

from . import _native
from .utils import _stream, compute_mu_lam_from_E_nu

N_HISTORY = 31   # 30 substeps per training step + the initial state (model.py:147-151, extra.py:199)


class GradField:
    """a field with an adjoint, both plain CUDA tensors"""

    def __init__(self, tensor: torch.Tensor, needs_grad=True):
        self.t = tensor
        self.grad = GradField(torch.zeros_like(tensor), needs_grad=False) if needs_grad else None

    @property
    def shape(self):
        return tuple(self.t.shape)

    def to_torch(self, device=None):
        return self.t if device is None else self.t.to(device)

    def to_numpy(self):
        return self.t.cpu().numpy()

    def from_torch(self, t):
        self.t.copy_(torch.as_tensor(t).to(self.t.device).reshape(self.t.shape))

    def fill(self, value):
        self.t.fill_(value)

    def copy_from(self, other):
        self.from_torch(other.to_torch())

    def ptr(self, s=None):
        """device pointer of the whole field or of history slice s"""
        return self.t.data_ptr() if s is None else self.t[s].data_ptr()


class MPM_state_opt:
    """model.py:136-223"""

    def __init__(self, n_particles, xyzs, covs, volumes, args, init_vel, device=None):
        dev = device or (xyzs.device if xyzs.is_cuda else torch.device("cuda", torch.cuda.current_device()))
        self.device = dev
        self.n_particles = n = int(n_particles)
        f32 = dict(dtype=torch.float32, device=dev)
        fit = bool(getattr(args, "fitting", True))
        self.particle_vol = GradField(torch.zeros(n, **f32), needs_grad=False)
        self.particle_density = GradField(torch.zeros(n, **f32), needs_grad=False)
        self.particle_mass = GradField(torch.zeros(n, **f32), needs_grad=False)
        self.particle_xyz = GradField(torch.zeros(N_HISTORY, n, 3, **f32), fit)
        self.particle_vel = GradField(torch.zeros(N_HISTORY, n, 3, **f32), fit)
        self.particle_F = GradField(torch.zeros(N_HISTORY, n, 3, 3, **f32), fit)
        self.particle_stress = GradField(torch.zeros(N_HISTORY, n, 3, 3, **f32), fit)
        self.particle_C = GradField(torch.zeros(N_HISTORY, n, 3, 3, **f32), fit)
        self.particle_init_cov = GradField(torch.zeros(n * 6, **f32), needs_grad=False)
        self.particle_cov = GradField(torch.zeros(n * 6, **f32), fit)

        self.particle_vol.from_torch(volumes.to(torch.float32))
        self.particle_density.fill(float(args.density))
        torch.mul(self.particle_density.t, self.particle_vol.t, out=self.particle_mass.t)   # utils.py:366-373
        if init_vel is None:
            init_vel = torch.zeros(n, 3)
        self.init2(xyzs, init_vel)
        self.init_identity_mat(self.particle_F)
        self.particle_init_cov.from_torch(covs.to(torch.float32).reshape(-1))
        self.particle_cov.copy_from(self.particle_init_cov)

        g = int(args.n_grid)
        self.grid_mass = GradField(torch.zeros(g, g, g, **f32), needs_grad=False)
        self.grid_v_in = GradField(torch.zeros(g, g, g, 3, **f32), fit)
        self.grid_v_out = GradField(torch.zeros(g, g, g, 3, **f32), fit)
        self.reset_grid_state()

    def init2(self, xyz_, vel_):          # model.py:177-185
        self.particle_xyz.t[0].copy_(torch.as_tensor(xyz_).to(self.device, torch.float32))
        self.particle_vel.t[0].copy_(torch.as_tensor(vel_).to(self.device, torch.float32))
        self.particle_stress.t[0].zero_()
        self.particle_C.t[0].zero_()

    def reset_grid_state(self):           # model.py:187-190
        self.grid_mass.t.zero_()
        self.grid_v_in.t.zero_()
        self.grid_v_out.t.zero_()

    def init_identity_mat(self, F):       # model.py:192-195
        F.t[0] = torch.eye(3, dtype=torch.float32, device=self.device)

    def set_grads(self, xyz_grad, cov_grad):   # model.py:197-208
        self.particle_xyz.grad.t[N_HISTORY - 1].copy_(torch.as_tensor(xyz_grad).to(self.device, torch.float32).reshape(self.n_particles, 3))
        self.particle_cov.grad.t.copy_(torch.as_tensor(cov_grad).to(self.device, torch.float32).reshape(-1))

    def clear_grad(self):                 # model.py:210-219
        for f in (self.particle_xyz, self.particle_vel, self.particle_F, self.particle_stress, self.particle_C, self.particle_cov,
                  self.grid_v_in, self.grid_v_out):
            f.grad.t.zero_()

    def cycle_init(self):                 # model.py:221-228
        last = N_HISTORY - 1
        for f in (self.particle_xyz, self.particle_vel, self.particle_F, self.particle_stress, self.particle_C):
            f.t[0].copy_(f.t[last])


class _FittingBackend:
    """the library context (grid geometry, error reporting, launch counting) without the packed particle arrays of
    the substep hot path: the fitting path keeps its own history-indexed state"""

    def __init__(self, n_grid, grid_extent, device):
        if not torch.cuda.is_available():
            raise _native.MpmkError("the fitting path needs a CUDA device; there is no CPU fallback")
        self.device = device
        idx = device.index if device.index is not None else torch.cuda.current_device()
        self.ctx = _native.Context(idx, int(n_grid), float(grid_extent), 1)


class FittingModel:
    """MPM_model with args.fitting (model.py:7-73): the parameters the differentiable kernels read, logE / y / mu / lam
    as original-order fields with adjoints"""

    def __init__(self, n_particles, args, device, ctx):
        import math
        from .utils import material_types
        self.args = args
        self.n_particles = n = int(n_particles)
        self.grid_extent = args.grid_extent
        self.n_grid = args.n_grid
        self.dx = args.grid_extent / args.n_grid
        self.inv_dx = args.n_grid / args.grid_extent
        if material_types.get(args.material, -1) < 0:
            raise TypeError("Material not supported yet")
        self.gravity = tuple(float(g) for g in args.gravity)
        f32 = dict(dtype=torch.float32, device=device)
        self.material = GradField(torch.full((n,), float(material_types[args.material]), **f32), needs_grad=False)
        self.logE = GradField(torch.full((n,), math.log10(args.E), **f32))
        self.y = GradField(torch.full((n,), -math.log(0.49 / args.nu - 1), **f32))
        self.mu = GradField(torch.zeros(n, **f32))
        self.lam = GradField(torch.zeros(n, **f32))
        self.viscosity = GradField(torch.full((n,), float(getattr(args, "viscosity", 0.05)), **f32), needs_grad=False)
        self._ctx = ctx
        compute_mu_lam_from_E_nu(ctx, n, self.logE.t, self.y.t, self.mu.t, self.lam.t)

    def clear_grad(self):                 # model.py:69-73
        for f in (self.logE, self.y, self.mu, self.lam):
            f.grad.t.zero_()


class FittingMixin:
    """the differentiable methods of MPM_Simulator (solver.py:54-108,131-133,168-177); `self.mpm_state` is an
    MPM_state_opt, `self.mpm_model` carries GradFields, `self._backend.ctx` the library context"""

    def _opt_call(self, name, *a):
        self._backend.ctx.call(name, *a, _stream())

    def _grid_box(self):
        """grid_postprocess[0] (solver.py:66,79,86): the one grid operation of the fitting path, a box boundary condition"""
        bc = self.grid_postprocess[0]
        if getattr(bc, "isCollide", False) or not hasattr(bc, "center"):
            raise NotImplementedError("the differentiable path applies a box boundary condition (set_bc_ground_only / fixed_cube) as grid_postprocess[0]")
        c = torch.tensor([float(v) for v in bc.center], dtype=torch.float32)
        s = torch.tensor([float(v) for v in bc.size], dtype=torch.float32)
        return c, s

    # strict_grad = True (default, like strict_q1 / strict_q3): the gradient Taichi's reverse mode produces for the
    # reference's kernel sequence, including where that is not the textbook one --
    #   * reset_grid_state (solver.py:73) resets values, not adjoints: grid_v_in.grad / grid_v_out.grad keep accumulating
    #     over the backward substeps of a frame (clear_grads is the only place that zeroes them);
    #   * compute_mu_lam_from_E_nu.grad runs in EVERY backward substep on the mu / lam adjoints accumulated so far (solver.py:90);
    #   * the boundary condition's constant store does not block the gradient of a stuck node.
    # strict_grad = False: the textbook adjoint of the forward pass (per-substep grid adjoints, one mu/lam chain-rule
    # step at the end, no flow through stuck nodes) -- what tests/test_gpu_fitting.py checks against float64 autograd.
    strict_grad = True

    def _forward_grid(self, dt, s):
        """p2g_opt + grid_normalization_and_gravity + grid_postprocess[0].apply on a freshly reset grid"""
        import ctypes as C
        st, md, n = self.mpm_state, self.mpm_model, self.n_particles
        st.reset_grid_state()
        self._opt_call("mpmk_opt_p2g", n, st.particle_xyz.ptr(s), st.particle_vel.ptr(s), st.particle_C.ptr(s), st.particle_stress.ptr(s),
                       st.particle_mass.ptr(), st.particle_vol.ptr(), st.grid_v_in.ptr(), st.grid_mass.ptr(), float(dt))
        c, sz = self._grid_box()
        g3 = (C.c_float * 3)(*[float(v) for v in md.gravity])
        c3 = (C.c_float * 3)(*c.tolist())
        s3 = (C.c_float * 3)(*sz.tolist())
        self._opt_call("mpmk_opt_grid", st.grid_v_in.ptr(), st.grid_mass.ptr(), st.grid_v_out.ptr(), float(dt), C.cast(g3, C.c_void_p),
                       C.cast(c3, C.c_void_p), C.cast(s3, C.c_void_p))

    def p2g2p_forward(self, dt, s):        # solver.py:54-70
        st, md, n = self.mpm_state, self.mpm_model, self.n_particles
        if not 0 <= s < N_HISTORY - 1:
            raise IndexError(f"substep index {s} outside the {N_HISTORY}-entry history")
        self._opt_call("mpmk_opt_stress", n, st.particle_F.ptr(s), md.mu.ptr(), md.lam.ptr(), st.particle_stress.ptr(s))
        self._forward_grid(dt, s)
        self._opt_call("mpmk_opt_g2p", n, st.particle_xyz.ptr(s), st.particle_F.ptr(s), st.grid_v_out.ptr(), st.particle_xyz.ptr(s + 1),
                       st.particle_vel.ptr(s + 1), st.particle_C.ptr(s + 1), st.particle_F.ptr(s + 1), float(dt))
        self.time += dt

    def p2g2p_backward(self, dt, s):       # solver.py:72-90
        st, md, n = self.mpm_state, self.mpm_model, self.n_particles
        # the grid history is not stored: redo p2g and the grid operations of substep s
        self._forward_grid(dt, s)
        self._opt_call("mpmk_opt_g2p_grad", n, st.particle_xyz.ptr(s), st.particle_F.ptr(s), st.grid_v_out.ptr(),
                       st.particle_xyz.grad.ptr(s + 1), st.particle_vel.grad.ptr(s + 1), st.particle_C.grad.ptr(s + 1), st.particle_F.grad.ptr(s + 1),
                       st.grid_v_out.grad.ptr(), st.particle_xyz.grad.ptr(s), st.particle_F.grad.ptr(s), float(dt))
        if self.strict_grad:
            self._opt_call("mpmk_opt_grid_grad", st.grid_mass.ptr(), st.grid_v_out.grad.ptr(), st.grid_v_in.grad.ptr(), None, None)
        else:
            import ctypes as C
            c, sz = self._grid_box()
            c3 = (C.c_float * 3)(*c.tolist())
            s3 = (C.c_float * 3)(*sz.tolist())
            self._opt_call("mpmk_opt_grid_grad", st.grid_mass.ptr(), st.grid_v_out.grad.ptr(), st.grid_v_in.grad.ptr(),
                           C.cast(c3, C.c_void_p), C.cast(s3, C.c_void_p))
        self._opt_call("mpmk_opt_p2g_grad", n, st.particle_xyz.ptr(s), st.particle_vel.ptr(s), st.particle_C.ptr(s), st.particle_stress.ptr(s),
                       st.particle_mass.ptr(), st.particle_vol.ptr(), st.grid_v_in.grad.ptr(), st.particle_xyz.grad.ptr(s),
                       st.particle_vel.grad.ptr(s), st.particle_C.grad.ptr(s), st.particle_stress.grad.ptr(s), float(dt))
        self._opt_call("mpmk_opt_stress_grad", n, st.particle_F.ptr(s), md.mu.ptr(), md.lam.ptr(), st.particle_stress.grad.ptr(s),
                       st.particle_F.grad.ptr(s), md.mu.grad.ptr(), md.lam.grad.ptr())
        if self.strict_grad or s == 0:
            self._opt_call("mpmk_opt_mu_lam_grad", n, md.logE.ptr(), md.y.ptr(), md.mu.grad.ptr(), md.lam.grad.ptr(), md.logE.grad.ptr(),
                           md.y.grad.ptr())
        if not self.strict_grad:
            st.grid_v_in.grad.t.zero_()
            st.grid_v_out.grad.t.zero_()

    def learn(self):                       # solver.py:92-108
        md = self.mpm_model
        self._opt_call("mpmk_opt_learn", self.n_particles, md.logE.ptr(), md.y.ptr(), md.logE.grad.ptr(), md.y.grad.ptr())

    def set_bc_ground_only(self):          # solver.py:131-133
        from .boundary_conditions import boundaryConditionTypeCallBacks
        self.grid_postprocess.append(boundaryConditionTypeCallBacks["sticky_ground"]())

    def postprocess_forward(self):         # solver.py:168-169
        st = self.mpm_state
        self._opt_call("mpmk_opt_cov", self.n_particles, st.particle_F.ptr(N_HISTORY - 1), st.particle_init_cov.ptr(), st.particle_cov.ptr())

    def postprocess_backward(self):        # solver.py:171-172
        st = self.mpm_state
        self._opt_call("mpmk_opt_cov_grad", self.n_particles, st.particle_F.ptr(N_HISTORY - 1), st.particle_init_cov.ptr(),
                       st.particle_cov.grad.ptr(), st.particle_F.grad.ptr(N_HISTORY - 1))

    def clear_grads(self):                 # solver.py:174-177
        self.mpm_model.clear_grad()
        self.mpm_state.clear_grad()

    def refresh_mu_lam(self):
        """compute_mu_lam_from_E_nu from the (learned) logE, y (the call extra.py:240 leaves commented out)"""
        md = self.mpm_model
        compute_mu_lam_from_E_nu(self._backend.ctx, self.n_particles, md.logE.t, md.y.t, md.mu.t, md.lam.t)
