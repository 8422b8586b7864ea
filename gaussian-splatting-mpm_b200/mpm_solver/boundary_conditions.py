"""Boundary conditions (reference: mpm_solver/boundary_conditions.py:7-117).

Same classes, same registries, same bc_args keys.  The grid box test (BasicBC.apply :23-27) and the
impulse (ImpulseBC.apply :41-45) run inside the fused grid / p2g kernels: the classes here carry the
parameters, evaluate the activity window on the host exactly as the reference does (:30-31, Python
floats) and turn themselves into `mpmk_grid_op` / `mpmk_impulse` records.  The init-time modifiers
(:48-85) are one-off masked assignments on the device.
"""
from __future__ import annotations

import math

import torch

from . import _native


class BasicBC:
    def __init__(self, n_particles, bc_args, sim_args):
        self.n_particles = n_particles
        self.substep_dt = sim_args.substep_dt

        self.id = bc_args["id"]
        self.type = bc_args["type"]

        self.start_time = bc_args["start_time"]
        self.end_time = bc_args["start_time"] + sim_args.substep_dt * bc_args["num_dt"]

        self.center = bc_args["center"]
        self.size = bc_args["size"]

        self.isCollide = False

    def isActive(self, time):
        return time >= self.start_time and time < self.end_time

    def as_grid_op(self) -> _native.GridOp:
        op = _native.GridOp()
        op.type = _native.OP_FIXED_CUBE
        op.a[:] = [float(x) for x in self.center]
        op.b[:] = [float(x) for x in self.size]
        op.c = 0.0
        return op

    def _inside(self, state):
        """all(|x_p - center| < size), in original particle order (bool[N])"""
        x = state._xyz_packed.to_torch()
        c = torch.tensor(self.center, dtype=torch.float32, device=x.device)
        s = torch.tensor(self.size, dtype=torch.float32, device=x.device)
        return ((x - c).abs() < s).all(dim=1)

    def apply(self, state, dx=None):
        raise NotImplementedError("fixed_cube runs inside the fused grid kernel (csrc/grid.cu); call p2g2p")


class ImpulseBC(BasicBC):
    def __init__(self, n_particles, bc_args, sim_args):
        self.force = [float(bc_args["force"][0]), float(bc_args["force"][1]), float(bc_args["force"][2])]
        super().__init__(n_particles, bc_args, sim_args)

    def as_impulse(self) -> _native.Impulse:
        im = _native.Impulse()
        im.center[:] = [float(x) for x in self.center]
        im.size[:] = [float(x) for x in self.size]
        im.force[:] = self.force
        im.substep_dt = float(self.substep_dt)
        return im

    def apply(self, state):
        raise NotImplementedError("impulse runs inside the p2g kernel prologue (csrc/p2g.cu); call p2g2p")


class MaterialParamsModifier(BasicBC):
    """additional_params.  The reference dereferences model.E / model.nu, which do not exist
    (boundary_conditions.py:57-70, solver.py:128): it crashes.  Intended behaviour implemented here
    (parity unpinned): per-particle E, nu, density override inside the box, mu/lam
    recomputed, then mu overridden unless it is the 1000 sentinel (:69)."""

    def __init__(self, n_particles, bc_args, sim_args):
        self.mu = bc_args["mu"]
        self.density = bc_args["density"]
        self.E = bc_args["E"]
        self.nu = bc_args["nu"]
        self.isMaterial = False
        super().__init__(n_particles, bc_args, sim_args)

    def apply(self, state, model):
        inside = self._inside(state)
        model._logE[inside] = math.log10(self.E)
        model._y[inside] = -math.log(0.49 / self.nu - 1)
        state.particle_density.t[inside] = float(self.density)

    def applymu(self, state, model):
        if self.mu != 1000:
            inside = self._inside(state)
            mu = model.mu.to_torch()
            mu[inside] = float(self.mu)
            model.mu.from_torch(mu)


class MaterialTypeModifier(BasicBC):
    def __init__(self, n_particles, bc_args, sim_args):
        self.material = bc_args["material"]
        self.isMaterial = True
        super().__init__(n_particles, bc_args, sim_args)

    def apply(self, state, model):  # boundary_conditions.py:81-85
        inside = self._inside(state)
        mat = model.material.to_torch()
        mat[inside] = float(self.material)
        model.material.from_torch(mat)


class StickyGroundBC(BasicBC):
    """boundary_conditions.py:88-94 skips BasicBC.__init__ (no `type`, no `isCollide`): only usable from
    the differentiable path (set_bc_ground_only, mpm_solver/fitting.py), which applies it as grid_postprocess[0]."""

    def __init__(self):
        self.center = [1.0, 0.6, 1.0]
        self.size = [1.0, 0.1, 1.0]

    def isActive(self, time):
        return True


preprocess_bc = (
    "impulse"
)

postprocess_bc = (
    "fixed_cube",
    "sticky_ground"
)

init_bc = (
    "additional_params",
    "modify_material"
)

boundaryConditionTypeCallBacks = {
    "fixed_cube": BasicBC,
    "impulse": ImpulseBC,
    "sticky_ground": StickyGroundBC,
    "additional_params": MaterialParamsModifier,
    "modify_material": MaterialTypeModifier
}
