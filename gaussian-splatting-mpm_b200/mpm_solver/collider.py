"""Half-space collider (reference: mpm_solver/collider.py:5-49).

The projection itself (collider.py:13-44) runs inside the fused grid kernel (csrc/grid.cu); this
class carries the parameters and turns them into an `mpmk_grid_op`.
"""
from __future__ import annotations

from . import _native


class MPM_Collider:
    def __init__(self, point, normal, friction):
        self.point = tuple(float(x) for x in point)
        self.normal = tuple(float(x) for x in normal)
        self.friction = float(friction)
        self.isCollide = True

    def as_grid_op(self) -> _native.GridOp:
        op = _native.GridOp()
        op.type = _native.OP_COLLIDER
        op.a[:] = self.point
        op.b[:] = self.normal
        op.c = self.friction
        return op

    def collide(self, time, dt, state, model):
        raise NotImplementedError("the collider runs inside the fused grid kernel (csrc/grid.cu); "
                                  "register it with add_surface_collider and call p2g2p")


collideTypeCallBacks = {
    "ground": MPM_Collider
}
