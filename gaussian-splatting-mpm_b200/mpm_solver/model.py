"""MPM_model / MPM_state on torch CUDA tensors (reference: mpm_solver/model.py:7-132).

The reference keeps every particle attribute in its own Taichi field.  Here the hot attributes live
in a packed float4 SoA (include/mpmk.h `mpmk_particle_set`) whose physical order is the cell-sorted
order of the last substep; `FieldView` presents each reference field (`particle_xyz`, `particle_F`,
`mu`, ...) in ORIGINAL particle order through `.to_torch()/.from_torch()/.fill()/.copy_from()`.
torch owns all particle and grid memory; the CUDA library borrows pointers.
"""
from __future__ import annotations

import ctypes as C
import math

import torch

from . import _native
from .utils import material_types, compute_mu_lam_from_E_nu, compute_mass_from_vol_density, _stream


def _require_cuda():
    if not torch.cuda.is_available():
        raise _native.MpmkError("mpm_solver needs a CUDA device (sm_100a); there is no CPU fallback")


class _Backend:
    """Packed device state + the libmpmk context shared by MPM_model and MPM_state."""

    F4 = ("xm", "vv", "c0", "c1", "cs", "s1", "f0", "f1", "f2", "cv0", "cv1")

    def __init__(self, n_particles: int, n_grid: int, grid_extent: float, device: torch.device, capacity=None,
                 ids=None, n_global=None):
        """capacity > n_particles leaves room for particles that migrate in (multi-GPU slabs, slabs.py);
        ids / n_global: global particle ids of the local particles and the global particle count -- the
        original-order arrays (init_cov, export buffers) are then indexed by global id."""
        _require_cuda()
        self.n = int(n_particles)
        self.capacity = int(capacity or n_particles)
        self.ids_init = None if ids is None else ids.to(device=device, dtype=torch.int32)
        self.n_global = int(n_global if n_global is not None else n_particles)
        self.n_grid = int(n_grid)
        self.device = device
        dev_index = device.index if device.index is not None else torch.cuda.current_device()
        self.ctx = _native.Context(dev_index, n_grid, grid_extent, self.capacity)
        cap = max(self.capacity, 1)
        z = lambda *shape, dtype=torch.float32: torch.zeros(*shape, dtype=dtype, device=device)
        self.sets = []
        for _ in range(2):
            d = {k: z(cap, 4) for k in self.F4}
            d["meta"] = z(cap, 4, dtype=torch.int32)
            self.sets.append(d)
        self.ft = [z(cap, 4) for _ in range(3)]
        self.keys = z(cap, dtype=torch.int32)
        g = self.n_grid
        self.grid_in = z(g, g, g, 4)
        self.grid_out = z(g, g, g, 4)
        # original-order, cold
        ng = max(self.n_global, 1)
        self.init_cov = z(ng, 6)
        self.out_xyz = z(ng, 3)
        self.out_cov = z(ng, 6)
        self.out_R = z(ng, 3, 3)
        self.params = _native.Params()
        self._bind()
        self.export_fresh = False     # out_* hold postprocess() results of the current state
        self._ftrial_valid = False
        self.cov_from_postprocess = False
        # The library evaluates compute_stress_from_F_trial for substep s+1 in the g2p epilogue of substep s; the
        # reference evaluates it at the start of substep s+1 (solver.py:36) with whatever was written to the model in
        # between.  Writes that the update depends on mark it stale; the next substep redoes it first (flush_stress).
        self.pristine = True          # F = I everywhere and no substep yet: the update is the identity, stress = 0
        self.stress_dirty = None      # None | {"use_ft": bool, "use_ys_before": bool}

    # -- plumbing -------------------------------------------------------------------------
    def _bind(self):
        b = _native.Bindings()
        for i in range(2):
            for k in _native.SET_FIELDS:
                setattr(b.set[i], k, self.sets[i][k].data_ptr())
        b.ft0, b.ft1, b.ft2 = (t.data_ptr() for t in self.ft)
        b.keys = self.keys.data_ptr()
        b.grid_in = self.grid_in.data_ptr()
        b.grid_out = self.grid_out.data_ptr()
        self.ctx.call("mpmk_bind", C.byref(b))
        self.ctx.call("mpmk_set_count", self.n, 0)

    def push_params(self):
        self.ctx.call("mpmk_set_params", C.byref(self.params))

    @property
    def cur(self):
        return self.sets[self.ctx.L.mpmk_current_set(self.ctx.h)]

    def ids(self):
        """original particle id of each physical slot (int64[n])"""
        return self.cur["meta"][: self.n, 0].long()

    def local_ids(self):
        """ids() for scattering into n-sized original-order tensors; slabs hold a subset of a global id
        space, where the reference's per-field original-order view does not exist"""
        if self.ids_init is not None:
            raise _native.MpmkError("original-order field views are not available on a slab; use SlabSimulator.gather()/export()")
        return self.ids()

    def reset_order(self):
        """physical order := original order, in set 0"""
        self.ctx.call("mpmk_set_count", self.n, 0)
        self.sets[0]["meta"][: self.n, 0] = (self.ids_init[: self.n] if self.ids_init is not None else
                                             torch.arange(self.n, dtype=torch.int32, device=self.device))
        self.sets[0]["meta"][: self.n, 3] = 0

    def invalidate(self):
        self.export_fresh = False
        self.cov_from_postprocess = False

    def read_small(self, t):
        """a small device tensor on the host without a copy engine (mpmk_read_small: a kernel stores it into pinned host
        memory).  `.item()` / `.cpu()` queue on the device-to-host copy engine behind whatever bulk download the caller
        has in flight on another stream -- a frame loop that streams its results out then stalls here for the rest of
        that download (measured: 3 of 4.3 ms per frame at 2 GPUs, 10 of 12.8 ms at 8)."""
        if t.dtype == torch.bool:
            t = t.to(torch.int32)
        t = t.contiguous()
        words = (t.numel() * t.element_size() + 3) // 4
        if getattr(self, "_pin", None) is None or self._pin.numel() < words:
            self._pin = torch.empty(max(words, 1024), dtype=torch.int32).pin_memory()
        self.ctx.call("mpmk_read_small", self._pin.data_ptr(), t.data_ptr(), words, _stream())
        return self._pin[:words].clone().view(t.dtype)[: t.numel()].reshape(t.shape)

    def mark_stress_dirty(self, f_written=False, ys_written=False):
        """a write the constitutive update depends on (mu, lam, material, yield_stress, F, the scalar model
        parameters): redo the update before the next substep, from the stored F_trial / pre-return-map yield stress
        where the library still has them, so that the result is what the reference computes at the start of p2g2p"""
        self.invalidate()
        if f_written:
            self.pristine = False
            if self._ftrial_valid:   # the stored F_trial describes a state the caller has just replaced
                self.ctx.call("mpmk_set_count", self.n, self.ctx.L.mpmk_current_set(self.ctx.h))
                self._ftrial_valid = False
        elif self.pristine:
            return
        d = self.stress_dirty or {"use_ft": True, "use_ys_before": True}
        if f_written:
            d["use_ft"] = False
            d["use_ys_before"] = False
        if ys_written:
            d["use_ys_before"] = False
        self.stress_dirty = d

    def flush_stress(self, dt):
        d = self.stress_dirty
        if d is not None:
            self.ctx.call("mpmk_constitutive_ex", float(dt), int(d["use_ft"]), int(d["use_ys_before"]), _stream())
            self.stress_dirty = None
        self.pristine = False

    def ftrial_valid(self):
        return self._ftrial_valid

    def track_cov(self):
        return bool(self.params.flags & _native.TRACK_COV)

    def recompute_keys(self):
        self.ctx.call("mpmk_compute_keys", _stream())

    def export(self, want_R=True):
        self.ctx.call("mpmk_export", self.init_cov.data_ptr(), self.out_xyz.data_ptr(), self.out_cov.data_ptr(),
                      self.out_R.data_ptr() if want_R else None, _stream())
        self.export_fresh = True


class FieldView:
    """One reference field, in original particle order, backed by columns of the packed arrays.

    cols: list of (array_name, column) giving the field's flattened components in order.
    """

    def __init__(self, backend: _Backend, cols, shape_tail, dtype=torch.float32, after_write=None):
        self._b = backend
        self._cols = cols
        self._tail = tuple(shape_tail)
        self._dtype = dtype
        self._after_write = after_write

    @property
    def shape(self):
        return (self._b.n,) + self._tail

    def _gather_sorted(self):
        cur, n = self._b.cur, self._b.n
        return torch.stack([cur[a][:n, c] for a, c in self._cols], dim=1)

    def to_torch(self, device=None):
        b = self._b
        vals = self._gather_sorted().to(self._dtype)
        out = torch.empty_like(vals)
        out[b.local_ids()] = vals
        out = out.reshape((b.n,) + self._tail)
        return out if device is None else out.to(device)

    def to_numpy(self):
        return self.to_torch().cpu().numpy()

    def from_torch(self, t):
        b = self._b
        t = t.to(device=b.device).reshape(b.n, -1)
        vals = t[b.local_ids()]
        cur = b.cur
        for k, (a, c) in enumerate(self._cols):
            cur[a][: b.n, c] = vals[:, k].to(cur[a].dtype)
        b.invalidate()
        if self._after_write:
            self._after_write()

    def fill(self, value):
        b = self._b
        cur = b.cur
        n_comp = len(self._cols)
        if isinstance(value, (int, float)):
            value = [value] * n_comp
        for k, (a, c) in enumerate(self._cols):
            cur[a][: b.n, c] = value[k]
        b.invalidate()
        if self._after_write:
            self._after_write()

    def copy_from(self, other):
        self.from_torch(other.to_torch())


class PlainField:
    """A field kept as a plain original-order torch tensor (cold attributes)."""

    def __init__(self, tensor, after_write=None):
        self.t = tensor
        self._after_write = after_write

    @property
    def shape(self):
        return tuple(self.t.shape)

    def to_torch(self, device=None):
        return self.t if device is None else self.t.to(device)

    def to_numpy(self):
        return self.t.cpu().numpy()

    def from_torch(self, t):
        self.t.copy_(t.to(self.t.device).reshape(self.t.shape))
        if self._after_write:
            self._after_write()

    def fill(self, value):
        self.t.fill_(value)
        if self._after_write:
            self._after_write()

    def copy_from(self, other):
        self.from_torch(other.to_torch())


class _MaterialField(FieldView):
    """model.material is an f32 field in the reference (model.py:25); stored as int32 in meta.y."""

    def to_torch(self, device=None):
        return super().to_torch(device).to(torch.float32)


class _VolField(FieldView):
    """particle_vol lives twice in the packed layout: s1.w (p2g) and meta.z bits (g2p)."""

    def _sync_bits(self):
        cur, n = self._b.cur, self._b.n
        cur["meta"][:n, 2] = cur["s1"][:n, 3].contiguous().view(torch.int32)

    def from_torch(self, t):
        super().from_torch(t)
        self._sync_bits()

    def fill(self, value):
        super().fill(value)
        self._sync_bits()


class _SymStressField(FieldView):
    """particle_stress [N,3,3] from the 6 stored symmetric components."""

    IDX = (0, 1, 2, 1, 3, 4, 2, 4, 5)

    def to_torch(self, device=None):
        b = self._b
        six = self._gather_sorted()
        full = six[:, list(self.IDX)]
        out = torch.empty_like(full)
        out[b.local_ids()] = full
        out = out.reshape(b.n, 3, 3)
        return out if device is None else out.to(device)

    def from_torch(self, t):
        t = t.to(self._b.device).reshape(self._b.n, 9)
        sym = 0.5 * (t + t.reshape(-1, 3, 3).transpose(1, 2).reshape(-1, 9))
        super().from_torch(sym[:, [0, 1, 2, 4, 5, 8]])

    def fill(self, value):
        super().fill(value)


class _FTrialField(FieldView):
    """particle_F_trial: the stored F_trial of the last g2p when present, else F (identical for
    elastic particles, and before the first substep)."""

    FT = [("ft0", 0), ("ft0", 1), ("ft0", 2), ("ft0", 3), ("ft1", 0), ("ft1", 1), ("ft1", 2), ("ft1", 3), ("ft2", 0)]

    def __init__(self, backend, f_cols):
        super().__init__(backend, f_cols, (3, 3))

    def _gather_sorted(self):
        b = self._b
        if b.ftrial_valid():
            n = b.n
            return torch.stack([b.ft[int(a[2])][:n, c] for a, c in self.FT], dim=1)
        return super()._gather_sorted()


class _YieldStressField(FieldView):
    """model.yield_stress as the reference holds it between substeps.  The library runs the return mapping of F_trial(s)
    in the epilogue of substep s (the reference: at the start of substep s+1, utils.py:14-54), so f2.w is one hardening /
    softening update ahead of the reference's field; the last g2p of a batch keeps the value from BEFORE that update next
    to F_trial (ft2.y), and that is what a reader gets while it is valid."""

    def _gather_sorted(self):
        b = self._b
        if b.ftrial_valid():
            return b.ft[2][: b.n, 1:2]
        return super()._gather_sorted()


class _ExportField:
    """particle_xyz / particle_cov / particle_R: produced by the export kernel straight into
    persistent torch buffers in original order (the zero-copy hand-off to the rasterizer that
    replaces main.py:311-312's host round trip)."""

    def __init__(self, backend, which, packed: FieldView | None):
        self._b = backend
        self._which = which
        self._packed = packed

    @property
    def shape(self):
        b = self._b
        return {"xyz": (b.n, 3), "cov": (b.n * 6,), "R": (b.n, 3, 3)}[self._which]

    def to_torch(self, device=None):
        b = self._b
        if self._which == "cov" and not b.cov_from_postprocess and b.track_cov():
            out = self._packed.to_torch().reshape(-1)      # update_cov's running value (utils.py:187-215)
        else:
            if not b.export_fresh:
                b.export()
            out = {"xyz": b.out_xyz, "cov": b.out_cov.view(-1), "R": b.out_R}[self._which]
        return out if device is None else out.to(device)

    def to_numpy(self):
        return self.to_torch().cpu().numpy()

    def from_torch(self, t):
        if self._packed is None:
            raise NotImplementedError(f"particle_{self._which} is an output of postprocess()")
        self._packed.from_torch(t.reshape(self._b.n, -1))

    def fill(self, value):
        if self._packed is None:
            raise NotImplementedError(f"particle_{self._which} is an output of postprocess()")
        self._packed.fill(value)

    def copy_from(self, other):
        self.from_torch(other.to_torch())


class _GridField:
    """Zero-copy view of the float4 node arrays in the reference's [n,n,n(,3)] layout (model.py:118-121)."""

    def __init__(self, tensor, sl):
        self._t = tensor
        self._sl = sl

    @property
    def shape(self):
        return tuple(self._t[..., self._sl].shape)

    def to_torch(self, device=None):
        v = self._t[..., self._sl]
        return v if device is None else v.to(device)

    def to_numpy(self):
        return self.to_torch().cpu().numpy()

    def fill(self, value):
        self._t[..., self._sl] = value


class MPM_model:
    """Per-particle material parameters + grid constants (reference: model.py:7-73)."""

    def __init__(self, n_particles: int, args, _backend: _Backend = None):
        self.args = args
        self.n_particles = n_particles
        # Eulerian grid settings, model.py:13-16
        self.grid_extent = args.grid_extent
        self.n_grid = args.n_grid
        self.dx = args.grid_extent / args.n_grid
        self.inv_dx = args.n_grid / args.grid_extent
        if _backend is None:
            _backend = _Backend(n_particles, args.n_grid, args.grid_extent, torch.device("cuda", torch.cuda.current_device()))
        self._backend = b = _backend
        self.init_general_params()
        self.init_elasiticity_params()
        self.init_plasticity_params()
        self.init_other_params()
        self._push()

    def init_general_params(self):  # model.py:24-33
        b = self._backend
        args_material = material_types.get(self.args.material, -1)
        if args_material < 0:
            raise TypeError("Material not supported yet")
        self.gravity = tuple(float(g) for g in self.args.gravity)
        self.material = _MaterialField(b, [("meta", 1)], (), dtype=torch.int32, after_write=b.mark_stress_dirty)
        self._args_material = args_material
        b.reset_order()
        b.sets[0]["meta"][: b.n, 1] = args_material

    def init_elasiticity_params(self):  # model.py:35-46
        b = self._backend
        n = max(b.n, 1)
        dev = b.device
        self._logE = torch.full((n,), math.log10(self.args.E), dtype=torch.float32, device=dev)
        self._y = torch.full((n,), -math.log(0.49 / self.args.nu - 1), dtype=torch.float32, device=dev)
        self.logE = PlainField(self._logE)
        self.y = PlainField(self._y)
        self.mu = FieldView(b, [("f2", 1)], (), after_write=b.mark_stress_dirty)
        self.lam = FieldView(b, [("f2", 2)], (), after_write=b.mark_stress_dirty)
        self.viscosity = PlainField(torch.full((n,), float(getattr(self.args, "viscosity", 0.05)), dtype=torch.float32, device=dev))
        self.recompute_mu_lam()

    def recompute_mu_lam(self):
        """compute_mu_lam_from_E_nu (utils.py:350-362) from logE, y into the packed mu, lam."""
        b = self._backend
        n = b.n
        if n == 0:
            return
        if b.ids_init is not None:
            # slab of a multi-GPU run: no per-particle E/nu modifiers there, the parameters are uniform
            mu = torch.empty(1, dtype=torch.float32, device=b.device)
            lam = torch.empty(1, dtype=torch.float32, device=b.device)
            compute_mu_lam_from_E_nu(b.ctx, 1, self._logE[:1].contiguous(), self._y[:1].contiguous(), mu, lam)
            self.mu.fill(mu)      # device scalars, broadcast on the device: no host round trip (a reload per frame would
            self.lam.fill(lam)    # queue two 4-byte downloads behind whatever bulk copy the caller has in flight)
            return
        mu = torch.empty(n, dtype=torch.float32, device=b.device)
        lam = torch.empty(n, dtype=torch.float32, device=b.device)
        compute_mu_lam_from_E_nu(b.ctx, n, self._logE, self._y, mu, lam)
        self.mu.from_torch(mu)
        self.lam.from_torch(lam)

    def init_plasticity_params(self):  # model.py:48-51
        self.friction_angle = 25.0
        sin_phi = math.sin(self.friction_angle / 180.0 * 3.141592653589793)
        self.alpha = math.sqrt(2.0 / 3.0) * 2.0 * sin_phi / (3.0 - sin_phi)

    def init_other_params(self):  # model.py:53-59
        b = self._backend
        self.yield_stress = _YieldStressField(b, [("f2", 3)], (), after_write=lambda: b.mark_stress_dirty(ys_written=True))
        self.yield_stress.fill(0.005)
        self.hardening = 1
        self.xi = 1
        self.plastic_viscosity = 0.008
        self.softening = 1.0

    def _reinit_packed(self):
        """refill the per-particle model parameters that live in the packed arrays (material, mu, lam,
        yield_stress) from the constructor arguments -- used when the particle set is reloaded."""
        b = self._backend
        b.cur["meta"][: b.n, 1] = self._args_material
        self._logE.fill_(math.log10(self.args.E))
        self._y.fill_(-math.log(0.49 / self.args.nu - 1))
        self.recompute_mu_lam()
        self.yield_stress.fill(0.005)

    def _push(self):
        """send the scalar model parameters to the library"""
        b = self._backend
        p = b.params
        p.gravity[:] = self.gravity
        p.alpha = self.alpha
        p.plastic_viscosity = self.plastic_viscosity
        p.xi = self.xi
        p.softening = self.softening
        p.hardening = int(self.hardening)
        b.push_params()

    def clear_grad(self):
        raise NotImplementedError("adjoint fields exist on the fitting model only (args.fitting = True: mpm_solver/fitting.py FittingModel)")


class MPM_state:
    """Particle + grid state (reference: model.py:77-132)."""

    def __init__(self, n_particles, xyzs, covs, volumes, args, _backend: _Backend = None):
        self.n_particles = n_particles
        if _backend is None:
            raise _native.MpmkError("MPM_state is constructed by MPM_Simulator (it shares the packed backend with MPM_model)")
        self._backend = b = _backend
        dev = b.device
        n = b.n
        F = [("f0", 0), ("f0", 1), ("f0", 2), ("f0", 3), ("f1", 0), ("f1", 1), ("f1", 2), ("f1", 3), ("f2", 0)]
        Cc = [("c0", 0), ("c0", 1), ("c0", 2), ("c0", 3), ("c1", 0), ("c1", 1), ("c1", 2), ("c1", 3), ("cs", 0)]
        S6 = [("cs", 1), ("cs", 2), ("cs", 3), ("s1", 0), ("s1", 1), ("s1", 2)]
        CV = [("cv0", 0), ("cv0", 1), ("cv0", 2), ("cv0", 3), ("cv1", 0), ("cv1", 1)]
        inval = b.invalidate
        self.particle_vol = _VolField(b, [("s1", 3)], (), after_write=inval)
        self.particle_density = PlainField(torch.full((max(n, 1),), float(args.density), dtype=torch.float32, device=dev))
        self.particle_mass = FieldView(b, [("xm", 3)], (), after_write=inval)
        self._xyz_packed = FieldView(b, [("xm", 0), ("xm", 1), ("xm", 2)], (3,), after_write=b.recompute_keys)
        self.particle_xyz = _ExportField(b, "xyz", self._xyz_packed)
        self.particle_vel = FieldView(b, [("vv", 0), ("vv", 1), ("vv", 2)], (3,), after_write=inval)
        self.particle_F = FieldView(b, F, (3, 3), after_write=lambda: b.mark_stress_dirty(f_written=True))
        # (the reference overwrites particle_stress from F_trial at the start of every p2g2p, utils.py:14-54)
        self.particle_stress = _SymStressField(b, S6, (3, 3), after_write=b.mark_stress_dirty)
        self.particle_C = FieldView(b, Cc, (3, 3), after_write=inval)
        self.particle_pressure = PlainField(torch.zeros(max(n, 1), dtype=torch.float32, device=dev))
        self.particle_init_cov = PlainField(b.init_cov.view(-1), after_write=inval)
        self._cov_packed = FieldView(b, CV, (6,), after_write=inval)
        self.particle_cov = _ExportField(b, "cov", self._cov_packed)
        self.particle_R = _ExportField(b, "R", None)
        self.particle_F_trial = _FTrialField(b, F)

        # initialisation, model.py:101-115
        self._args_density = float(args.density)
        self._init_particles(xyzs, covs, volumes)

        # grid, model.py:117-122 -- zero-copy views of the float4 node arrays
        self.grid_mass = _GridField(b.grid_in, 3)
        self.grid_v_in = _GridField(b.grid_in, slice(0, 3))
        self.grid_v_out = _GridField(b.grid_out, slice(0, 3))

    def _init_particles(self, xyzs, covs, volumes):
        """model.py:101-115 on the packed arrays; physical order == original order afterwards."""
        b = self._backend
        n, dev = b.n, b.device
        if n == 0:
            return
        b.reset_order()
        cur = b.cur
        vol = volumes.to(device=dev, dtype=torch.float32, non_blocking=True).reshape(n)
        cur["s1"][:n, 3] = vol
        cur["meta"][:n, 2] = vol.contiguous().view(torch.int32)
        compute_mass_from_vol_density(n, self.particle_density.t, vol, cur["xm"][:n, 3])
        cur["xm"][:n, :3] = xyzs.to(device=dev, dtype=torch.float32, non_blocking=True).reshape(n, 3)
        cur["vv"][:n].zero_()
        for a in ("c0", "c1", "cs"):
            cur[a][:n].zero_()
        cur["s1"][:n, :3] = 0.0
        for a in ("f0", "f1"):     # F = I: (1, 0, 0, 0), (1, 0, 0, 0), (1, ...)
            cur[a][:n].zero_()
            cur[a][:n, 0] = 1.0
        cur["f2"][:n, 0] = 1.0
        c6 = covs.to(device=dev, dtype=torch.float32, non_blocking=True).reshape(n, 6)
        if b.ids_init is not None:
            b.init_cov[b.ids_init[:n].long()] = c6
        else:
            b.init_cov.copy_(c6)
        cur["cv0"][:n] = c6[:, :4]
        cur["cv1"][:n, :2] = c6[:, 4:]
        b._ftrial_valid = False
        b.pristine = True
        b.stress_dirty = None
        b.invalidate()
        b.recompute_keys()

    def reset_grid_state(self):
        """model.py:124-127.  The substep clears only the blocks it touched; this is the dense version."""
        self._backend.grid_in.zero_()
        self._backend.grid_out.zero_()

    def init_identity_mat(self, F):  # model.py:129-132
        F.fill([1.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 1.0])



class MPM_state:
    """Particle + grid state (reference: model.py:77-132)."""

    def __init__(self, n_particles, xyzs, covs, volumes, args, _backend: _Backend = None):
        self.n_particles = n_particles
        if _backend is None:
            raise _native.MpmkError("MPM_state is constructed by MPM_Simulator (it shares the packed backend with MPM_model)")
        self._backend = b = _backend
        dev = b.device
        n = b.n
        F = [("f0", 0), ("f0", 1), ("f0", 2), ("f0", 3), ("f1", 0), ("f1", 1), ("f1", 2), ("f1", 3), ("f2", 0)]
        Cc = [("c0", 0), ("c0", 1), ("c0", 2), ("c0", 3), ("c1", 0), ("c1", 1), ("c1", 2), ("c1", 3), ("cs", 0)]
        S6 = [("cs", 1), ("cs", 2), ("cs", 3), ("s1", 0), ("s1", 1), ("s1", 2)]
        CV = [("cv0", 0), ("cv0", 1), ("cv0", 2), ("cv0", 3), ("cv1", 0), ("cv1", 1)]
        inval = b.invalidate
        self.particle_vol = _VolField(b, [("s1", 3)], (), after_write=inval)
        self.particle_density = PlainField(torch.full((max(n, 1),), float(args.density), dtype=torch.float32, device=dev))
        self.particle_mass = FieldView(b, [("xm", 3)], (), after_write=inval)
        self._xyz_packed = FieldView(b, [("xm", 0), ("xm", 1), ("xm", 2)], (3,), after_write=b.recompute_keys)
        self.particle_xyz = _ExportField(b, "xyz", self._xyz_packed)
        self.particle_vel = FieldView(b, [("vv", 0), ("vv", 1), ("vv", 2)], (3,), after_write=inval)
        self.particle_F = FieldView(b, F, (3, 3), after_write=lambda: b.mark_stress_dirty(f_written=True))
        # (the reference overwrites particle_stress from F_trial at the start of every p2g2p, utils.py:14-54)
        self.particle_stress = _SymStressField(b, S6, (3, 3), after_write=b.mark_stress_dirty)
        self.particle_C = FieldView(b, Cc, (3, 3), after_write=inval)
        self.particle_pressure = PlainField(torch.zeros(max(n, 1), dtype=torch.float32, device=dev))
        self.particle_init_cov = PlainField(b.init_cov.view(-1), after_write=inval)
        self._cov_packed = FieldView(b, CV, (6,), after_write=inval)
        self.particle_cov = _ExportField(b, "cov", self._cov_packed)
        self.particle_R = _ExportField(b, "R", None)
        self.particle_F_trial = _FTrialField(b, F)

        # initialisation, model.py:101-115
        self._args_density = float(args.density)
        self._init_particles(xyzs, covs, volumes)

        # grid, model.py:117-122 -- zero-copy views of the float4 node arrays
        self.grid_mass = _GridField(b.grid_in, 3)
        self.grid_v_in = _GridField(b.grid_in, slice(0, 3))
        self.grid_v_out = _GridField(b.grid_out, slice(0, 3))

    def _init_particles(self, xyzs, covs, volumes):
        """model.py:101-115 on the packed arrays; physical order == original order afterwards."""
        b = self._backend
        n, dev = b.n, b.device
        if n == 0:
            return
        b.reset_order()
        cur = b.cur
        vol = volumes.to(device=dev, dtype=torch.float32, non_blocking=True).reshape(n)
        cur["s1"][:n, 3] = vol
        cur["meta"][:n, 2] = vol.contiguous().view(torch.int32)
        compute_mass_from_vol_density(n, self.particle_density.t, vol, cur["xm"][:n, 3])
        cur["xm"][:n, :3] = xyzs.to(device=dev, dtype=torch.float32, non_blocking=True).reshape(n, 3)
        cur["vv"][:n].zero_()
        for a in ("c0", "c1", "cs"):
            cur[a][:n].zero_()
        cur["s1"][:n, :3] = 0.0
        for a in ("f0", "f1"):     # F = I: (1, 0, 0, 0), (1, 0, 0, 0), (1, ...)
            cur[a][:n].zero_()
            cur[a][:n, 0] = 1.0
        cur["f2"][:n, 0] = 1.0
        c6 = covs.to(device=dev, dtype=torch.float32, non_blocking=True).reshape(n, 6)
        if b.ids_init is not None:
            b.init_cov[b.ids_init[:n].long()] = c6
        else:
            b.init_cov.copy_(c6)
        cur["cv0"][:n] = c6[:, :4]
        cur["cv1"][:n, :2] = c6[:, 4:]
        b._ftrial_valid = False
        b.pristine = True
        b.stress_dirty = None
        b.invalidate()
        b.recompute_keys()

    def reset_grid_state(self):
        """model.py:124-127.  The substep clears only the blocks it touched; this is the dense version."""
        self._backend.grid_in.zero_()
        self._backend.grid_out.zero_()

    def init_identity_mat(self, F):  # model.py:129-132
        F.fill([1.0, 0.0, 0.0, 0.0, 1.0, 0.0, 0.0, 0.0, 1.0])
