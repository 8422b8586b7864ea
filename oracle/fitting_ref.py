"""TEST INFRASTRUCTURE ONLY (see oracle/mpm_oracle.c's header): float64 CPU restatement of the reference's
differentiable path, checked against by tests/test_gpu_fitting.py.  Never imported by the product.

Each forward kernel of the reference is restated as a differentiable torch function (float64):
    stress_fn   mpm_solver/utils.py:56-76    compute_stress_from_F_opt
    p2g_fn      mpm_solver/utils.py:137-174  p2g_opt
    norm_fn     mpm_solver/utils.py:178-183  grid_normalization_and_gravity
    box_fn      mpm_solver/boundary_conditions.py:23-27  BasicBC.apply / StickyGroundBC
    g2p_fn      mpm_solver/utils.py:285-347  g2p_opt
    cov_fn      mpm_solver/utils.py:436-467  compute_cov_from_F_opt
    mu_lam_fn   mpm_solver/utils.py:350-362  compute_mu_lam_from_E_nu
and the backward sweep is orchestrated the way solver.py:72-90 / extra.py:222-233 call Taichi's per-kernel `.grad`
kernels on PERSISTENT, ACCUMULATING adjoint fields -- the vector-Jacobian product of each kernel comes from
torch.autograd on that kernel alone.  `strict=True` reproduces Taichi's reverse mode literally (grid adjoints are never
reset between backward substeps, compute_mu_lam_from_E_nu.grad runs every substep on the accumulated mu / lam adjoints,
the boundary condition's constant store does not block a stuck node's gradient); `strict=False` is the textbook
gradient, which is also what end-to-end autograd through the whole forward pass gives (checked in the test).

Parity unpinned: Taichi cannot run here, so the `strict` orchestration restates Taichi 1.5's documented reverse-mode
semantics (global-store adjoints are read, never cleared), not its output."""
import math

import torch

F64 = torch.float64


def weights(x, inv_dx, n_grid):
    gp = x * inv_dx
    base = (gp - 0.5).to(torch.int64).clamp(0, n_grid - 3)       # cast: no gradient (and this build's Q9 clamp)
    fx = gp - base.to(F64)
    w = torch.stack([0.5 * (1.5 - fx) ** 2, 0.75 - (fx - 1.0) ** 2, 0.5 * (fx - 0.5) ** 2], -1)     # [N,3(axis),3(node)]
    dw = torch.stack([fx - 1.5, -2.0 * (fx - 1.0), fx - 0.5], -1)
    return base, fx, w, dw


def _nodes(base, n_grid):
    off = torch.tensor([[i, j, k] for i in range(3) for j in range(3) for k in range(3)], dtype=torch.int64)   # [27,3]
    idx = base[:, None, :] + off[None]
    return off, (idx[..., 0] * n_grid + idx[..., 1]) * n_grid + idx[..., 2]       # [N,27]


def _w27(w, dw, off, inv_dx):
    wx, wy, wz = w[:, 0][:, off[:, 0]], w[:, 1][:, off[:, 1]], w[:, 2][:, off[:, 2]]
    dx_, dy_, dz_ = dw[:, 0][:, off[:, 0]], dw[:, 1][:, off[:, 1]], dw[:, 2][:, off[:, 2]]
    weight = wx * wy * wz
    dweight = torch.stack([dx_ * wy * wz, wx * dy_ * wz, wx * wy * dz_], -1) * inv_dx      # [N,27,3]
    return weight, dweight


def stress_fn(F, mu, lam):
    J = torch.linalg.det(F)
    small = J.abs() < 1e-2
    Jc = torch.where(small, (1e-2 * torch.sign(J)).detach(), J)
    I = torch.eye(3, dtype=F64)
    E = 0.5 * (F.transpose(1, 2) @ F - I)
    S = 2 * mu[:, None, None] * E + lam[:, None, None] * E.diagonal(dim1=1, dim2=2).sum(-1)[:, None, None] * I
    return F @ S @ F.transpose(1, 2) / Jc[:, None, None]


def p2g_fn(x, v, C, stress, mass, vol, n_grid, dx, inv_dx, dt, detach_mass=True):
    """detach_mass: grid_mass is a field without needs_grad in the reference (model.py:172), so nothing flows back through the
    dependence of the node masses on the particle positions; False keeps that term (the true derivative, for the
    finite-difference check of this file's own autograd plumbing)"""
    base, fx, w, dw = weights(x, inv_dx, n_grid)
    off, node = _nodes(base, n_grid)
    weight, dweight = _w27(w, dw, off, inv_dx)
    dpos = (off[None].to(F64) - fx[:, None, :]) * dx                                   # [N,27,3]
    aff = v[:, None, :] + torch.einsum("ncd,nkd->nkc", C, dpos)
    ef = -vol[:, None, None] * torch.einsum("ncd,nkd->nkc", stress, dweight)
    add = weight[..., None] * mass[:, None, None] * aff + dt * ef
    gv = torch.zeros(n_grid ** 3, 3, dtype=F64).index_add(0, node.reshape(-1), add.reshape(-1, 3))
    wm = (weight * mass[:, None]).reshape(-1)
    gm = torch.zeros(n_grid ** 3, dtype=F64).index_add(0, node.reshape(-1), wm.detach() if detach_mass else wm)
    return gv, gm


def norm_fn(gv_in, gmass, dt, gravity):
    on = gmass > 1e-15
    safe = torch.where(on, gmass, torch.ones_like(gmass))
    return torch.where(on[:, None], gv_in / safe[:, None] + dt * gravity[None], torch.zeros_like(gv_in))


def box_mask(n_grid, dx, center, size):
    r = torch.arange(n_grid, dtype=torch.float32) * torch.tensor(dx, dtype=torch.float32)     # fp32 predicate, as the kernels
    c = [torch.tensor(v, dtype=torch.float32) for v in center]
    s = [torch.tensor(v, dtype=torch.float32) for v in size]
    mx, my, mz = ((r - c[d]).abs() < s[d] for d in range(3))
    return (mx[:, None, None] & my[None, :, None] & mz[None, None, :]).reshape(-1)


def box_fn(gv_out, mask):
    return torch.where(mask[:, None], torch.zeros_like(gv_out), gv_out)


def g2p_fn(x, F, gv_out, n_grid, inv_dx, dt):
    base, fx, w, dw = weights(x, inv_dx, n_grid)
    off, node = _nodes(base, n_grid)
    weight, dweight = _w27(w, dw, off, inv_dx)
    g = gv_out[node]                                                                    # [N,27,3]
    dpos = off[None].to(F64) - fx[:, None, :]
    new_v = (g * weight[..., None]).sum(1)
    new_C = torch.einsum("nkc,nkd->ncd", g * (weight * inv_dx * 4.0)[..., None], dpos)
    new_F = torch.einsum("nkc,nkd->ncd", g, dweight)
    I = torch.eye(3, dtype=F64)
    return x + dt * new_v, new_v, new_C, (I + dt * new_F) @ F


def _sym6(a):
    return torch.stack([torch.stack([a[:, 0], a[:, 1], a[:, 2]], -1), torch.stack([a[:, 1], a[:, 3], a[:, 4]], -1),
                        torch.stack([a[:, 2], a[:, 4], a[:, 5]], -1)], 1)


def cov_fn(F, init_cov6):
    c = F @ _sym6(init_cov6) @ F.transpose(1, 2)
    return torch.stack([c[:, 0, 0], c[:, 0, 1], c[:, 0, 2], c[:, 1, 1], c[:, 1, 2], c[:, 2, 2]], -1)


def mu_lam_fn(logE, y):
    E = 10.0 ** logE
    nu = 0.49 / (1.0 + torch.exp(-y))
    return E / (2.0 * (1.0 + nu)), E * nu / ((1.0 + nu) * (1.0 - 2.0 * nu))


class FittingRef:
    def __init__(self, xyz, cov6, vol, vel, n_grid, grid_extent, E, nu, density, gravity, box=None, n_steps=30):
        t = lambda a: torch.as_tensor(a, dtype=F64)
        self.n = len(xyz)
        self.n_grid, self.dx, self.inv_dx = n_grid, grid_extent / n_grid, n_grid / grid_extent
        self.S = n_steps
        self.x0, self.v0, self.init_cov, self.vol = t(xyz), t(vel), t(cov6).reshape(-1, 6), t(vol)
        self.mass = density * self.vol
        self.gravity = t(gravity)
        self.logE = torch.full((self.n,), math.log10(E), dtype=F64)
        self.y = torch.full((self.n,), -math.log(0.49 / nu - 1), dtype=F64)
        self.mask = box_mask(n_grid, float(torch.tensor(self.dx, dtype=torch.float32)), *box) if box else torch.zeros(n_grid ** 3, dtype=torch.bool)

    def forward(self, dt, logE=None, y=None, x0=None, detach_mass=True):
        """the whole forward pass as one differentiable function; returns the history"""
        logE = self.logE if logE is None else logE
        y = self.y if y is None else y
        mu, lam = mu_lam_fn(logE, y)
        x = [self.x0 if x0 is None else x0]; v = [self.v0]
        C = [torch.zeros(self.n, 3, 3, dtype=F64)]; F = [torch.eye(3, dtype=F64).repeat(self.n, 1, 1)]
        stress = []
        for s in range(self.S):
            stress.append(stress_fn(F[s], mu, lam))
            gv, gm = p2g_fn(x[s], v[s], C[s], stress[s], self.mass, self.vol, self.n_grid, self.dx, self.inv_dx, dt, detach_mass)
            gout = box_fn(norm_fn(gv, gm, dt, self.gravity), self.mask)
            x1, v1, C1, F1 = g2p_fn(x[s], F[s], gout, self.n_grid, self.inv_dx, dt)
            x.append(x1); v.append(v1); C.append(C1); F.append(F1)
        return dict(x=x, v=v, C=C, F=F, stress=stress, cov=cov_fn(F[-1], self.init_cov))

    def backward(self, dt, g_x_last, g_cov, strict):
        """solver.py:72-90 + extra.py:222-233 on persistent adjoint fields; returns (logE.grad, y.grad, xyz.grad[0])"""
        with torch.no_grad():
            H = self.forward(dt)
        S, n = self.S, self.n
        z = lambda *sh: torch.zeros(*sh, dtype=F64)
        gx = [z(n, 3) for _ in range(S + 1)]; gv = [z(n, 3) for _ in range(S + 1)]
        gC = [z(n, 3, 3) for _ in range(S + 1)]; gF = [z(n, 3, 3) for _ in range(S + 1)]
        gstress = [z(n, 3, 3) for _ in range(S)]
        g_gin, g_gout = z(self.n_grid ** 3, 3), z(self.n_grid ** 3, 3)
        g_mu, g_lam, g_logE, g_y = z(n), z(n), z(n), z(n)
        gx[S] += torch.as_tensor(g_x_last, dtype=F64)
        # postprocess_backward
        Fl = H["F"][S].clone().requires_grad_(True)
        gF[S] += torch.autograd.grad(cov_fn(Fl, self.init_cov), Fl, torch.as_tensor(g_cov, dtype=F64).reshape(-1, 6))[0]
        mu, lam = mu_lam_fn(self.logE, self.y)
        for s in reversed(range(S)):
            # redo p2g + grid of substep s
            with torch.no_grad():
                gvin, gm = p2g_fn(H["x"][s], H["v"][s], H["C"][s], H["stress"][s], self.mass, self.vol, self.n_grid, self.dx, self.inv_dx, dt)
                gout = box_fn(norm_fn(gvin, gm, dt, self.gravity), self.mask)
            # g2p_opt.grad
            xs, Fs, go = (t.clone().requires_grad_(True) for t in (H["x"][s], H["F"][s], gout))
            outs = g2p_fn(xs, Fs, go, self.n_grid, self.inv_dx, dt)
            a, b, c = torch.autograd.grad(outs, (xs, Fs, go), (gx[s + 1], gv[s + 1], gC[s + 1], gF[s + 1]))
            gx[s] += a; gF[s] += b; g_gout += c
            # grid_normalization_and_gravity.grad (+ the boundary condition's, which Taichi leaves without effect)
            gi = gvin.clone().requires_grad_(True)
            out = norm_fn(gi, gm, dt, self.gravity)
            if not strict:
                out = box_fn(out, self.mask)
            g_gin += torch.autograd.grad(out, gi, g_gout)[0]
            # p2g_opt.grad
            xs, vs, Cs, ss = (t.clone().requires_grad_(True) for t in (H["x"][s], H["v"][s], H["C"][s], H["stress"][s]))
            gvin2, _ = p2g_fn(xs, vs, Cs, ss, self.mass, self.vol, self.n_grid, self.dx, self.inv_dx, dt)
            a, b, c, d = torch.autograd.grad(gvin2, (xs, vs, Cs, ss), g_gin)
            gx[s] += a; gv[s] += b; gC[s] += c; gstress[s] += d
            # compute_stress_from_F_opt.grad
            Fs, m_, l_ = (t.clone().requires_grad_(True) for t in (H["F"][s], mu, lam))
            a, b, c = torch.autograd.grad(stress_fn(Fs, m_, l_), (Fs, m_, l_), gstress[s])
            gF[s] += a; g_mu += b; g_lam += c
            # compute_mu_lam_from_E_nu.grad
            if strict or s == 0:
                le, yy = self.logE.clone().requires_grad_(True), self.y.clone().requires_grad_(True)
                a, b = torch.autograd.grad(mu_lam_fn(le, yy), (le, yy), (g_mu, g_lam))
                g_logE += a; g_y += b
            if not strict:
                g_gin.zero_(); g_gout.zero_()
        return g_logE, g_y, gx[0], dict(g_mu=g_mu, g_lam=g_lam, gF0=gF[0])
