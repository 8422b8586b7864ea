"""Deterministic synthetic particle clouds.

A block of grid cells filled with `ppc` stratified-jittered particles per cell (for ppc=8: one per
octant, uniform inside the octant), rigid spin about the block centre, random anisotropic
covariances.  numpy only, so the same cloud feeds the CUDA solver and the CPU oracle.
"""
from __future__ import annotations

import types

import numpy as np

SEED = 20240501

WORKLOADS = {
    # name: (n_grid, grid_extent, cells_lo, cells_n, dt)
    "S-4M": (256, 2.0, (88, 88, 88), (80, 80, 80), 1e-4),            # N = 4,096,000
    "S-32M": (512, 2.0, (176, 176, 176), (160, 160, 160), 5e-5),      # N = 32,768,000
    "S-32M-bar": (512, 2.0, (32, 208, 208), (448, 96, 96), 5e-5),     # N = 33,030,144 (multi-GPU slabs)
    "S-bar1": (512, 2.0, (32, 208, 208), (56, 96, 96), 5e-5),         # N = 4,128,768: one rank's share of S-32M-bar
    "S-512K": (128, 2.0, (44, 44, 44), (40, 40, 40), 2e-4),           # N = 512,000 (tests)
    "S-64K": (64, 2.0, (22, 22, 22), (20, 20, 20), 4e-4),             # N = 64,000 (tests)
}


def make_cloud(n_grid, grid_extent, cells_lo, cells_n, ppc=8, seed=SEED, omega=(0.0, 0.0, 2.0)):
    """Returns dict(xyz[N,3], vel[N,3], cov[N,6], vol[N]) float32, particle-major over cells (x slowest)."""
    rng = np.random.default_rng(seed)
    dx = grid_extent / n_grid
    cx, cy, cz = cells_n
    ncell = cx * cy * cz
    ii, jj, kk = np.meshgrid(np.arange(cx), np.arange(cy), np.arange(cz), indexing="ij")
    cell = np.stack([ii.ravel(), jj.ravel(), kk.ravel()], axis=1).astype(np.float64) + np.asarray(cells_lo, np.float64)
    if ppc == 8:
        octant = np.array([[a, b, c] for a in (0, 1) for b in (0, 1) for c in (0, 1)], np.float64) * 0.5
        jit = rng.random((ncell, 8, 3)) * 0.5
        frac = octant[None, :, :] + jit
    else:
        frac = rng.random((ncell, ppc, 3))
    # keep a small margin from cell faces so fp32 rounding cannot move a particle across a face at t=0
    frac = np.clip(frac, 1e-3, 1.0 - 1e-3)
    xyz = ((cell[:, None, :] + frac) * dx).reshape(-1, 3)
    n = xyz.shape[0]
    centre = (np.asarray(cells_lo, np.float64) + 0.5 * np.asarray(cells_n, np.float64)) * dx
    r = xyz - centre
    w = np.asarray(omega, np.float64)
    vel = np.cross(np.broadcast_to(w, r.shape), r)
    # covariance Q diag(s^2) Q^T, s ~ U(0.1, 0.5) dx, Q a random rotation (from a random unit quaternion)
    q = rng.standard_normal((n, 4))
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    a, b, c, d = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    Q = np.stack([
        np.stack([a * a + b * b - c * c - d * d, 2 * (b * c - a * d), 2 * (b * d + a * c)], 1),
        np.stack([2 * (b * c + a * d), a * a - b * b + c * c - d * d, 2 * (c * d - a * b)], 1),
        np.stack([2 * (b * d - a * c), 2 * (c * d + a * b), a * a - b * b - c * c + d * d], 1)], 1)
    s2 = (rng.uniform(0.1, 0.5, (n, 3)) * dx) ** 2
    S = np.einsum("nij,nj,nkj->nik", Q, s2, Q)
    cov = np.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], 1)
    vol = np.full(n, dx ** 3 / ppc)
    f = lambda x: np.ascontiguousarray(x, dtype=np.float32)
    return {"xyz": f(xyz), "vel": f(vel), "cov": f(cov), "vol": f(vol)}


BAR = dict(n_grid=512, extent=2.0, lo=(32, 208, 208), block=(56, 96, 96), dt=5e-5)   # S-32M-bar = 8 blocks along x


def bar_block(rank, world, seed=SEED):
    """rank's 56 x 96 x 96-cell block (4,128,768 particles) of the bar that `world` ranks simulate together
    (bench.py --gpus N, weak scaling); the rigid spin is about the centre of the WHOLE bar"""
    n_grid, extent, (bx0, by0, bz0), (wx, wy, wz) = BAR["n_grid"], BAR["extent"], BAR["lo"], BAR["block"]
    dx = extent / n_grid
    cloud = make_cloud(n_grid, extent, (bx0 + wx * rank, by0, bz0), (wx, wy, wz), seed=seed + rank)
    centre = np.array([(bx0 + 0.5 * wx * world) * dx, (by0 + 0.5 * wy) * dx, (bz0 + 0.5 * wz) * dx])
    r = cloud["xyz"].astype(np.float64) - centre
    cloud["vel"] = np.cross(np.broadcast_to(np.array([0.0, 0.0, 2.0]), r.shape), r).astype(np.float32)
    return cloud


def bar_workload(world, material="jelly"):
    """(args, ops, bounds) of the bar for `world` ranks"""
    n_grid, extent, (bx0, by0, bz0), (wx, wy, wz) = BAR["n_grid"], BAR["extent"], BAR["lo"], BAR["block"]
    args = sim_args(n_grid=n_grid, grid_extent=extent, substep_dt=BAR["dt"], material=material)
    ops = block_ops(n_grid, extent, (bx0, by0, bz0), (wx * world, wy, wz))
    bounds = [0] + [bx0 + wx * k for k in range(1, world)] + [n_grid]
    return args, ops, bounds


def bar_cloud(world, seed=SEED):
    """the whole bar of `world` blocks as one cloud (the reference arm of bench.py --gpus N runs it on the host cores)"""
    parts = [bar_block(r, world, seed) for r in range(world)]
    return {k: np.concatenate([p[k] for p in parts]) for k in parts[0]}


def workload(name, ppc=8, seed=SEED):
    n_grid, extent, lo, cn, dt = WORKLOADS[name]
    cloud = make_cloud(n_grid, extent, lo, cn, ppc=ppc, seed=seed)
    return cloud, sim_args(n_grid=n_grid, grid_extent=extent, substep_dt=dt), block_ops(n_grid, extent, lo, cn)


def sim_args(n_grid=64, grid_extent=2.0, material="jelly", E=2e5, nu=0.3, density=1000.0,
             gravity=(0.0, 0.0, -9.8), substep_dt=1e-4, frame_dt=1e-2, viscosity=0.05, **kw):
    """Namespace with the MPMParams fields the solver reads (arguments/__init__.py:46-85)."""
    ns = types.SimpleNamespace(n_grid=n_grid, grid_extent=grid_extent, material=material, E=E, nu=nu,
                               density=density, gravity=list(gravity), substep_dt=substep_dt, frame_dt=frame_dt,
                               viscosity=viscosity, fitting=False, boundary_conditions=[])
    ns.steps_per_frame = int(frame_dt / substep_dt)
    for k, v in kw.items():
        setattr(ns, k, v)
    return ns


def block_ops(n_grid, grid_extent, cells_lo, cells_n):
    """One always-on fixed_cube slab under the block plus the z-plane collider main.py:276 always adds,
    placed relative to the block so the grid kernel's op list is exercised."""
    dx = grid_extent / n_grid
    z0 = cells_lo[2] * dx
    bcs = [{"id": 0, "type": "fixed_cube", "center": [grid_extent / 2, grid_extent / 2, z0 - 4 * dx],
            "size": [grid_extent, grid_extent, 2 * dx], "start_time": 0.0, "num_dt": 10 ** 12}]
    collider = {"point": (0.0, 0.0, z0 - 1.0 * dx), "normal": (0.0, 0.0, 1.0), "friction": 0.0}
    return {"boundary_conditions": bcs, "collider": collider}


# ---- config-shaped clouds --------------------------------------------------------------
# The reference's scenes are git-LFS pointers (no .ply payload in the tree), so the named configs are
# reproduced by SHAPE: particle count, grid, dt, material, gravity and the boundary-condition list of the
# JSON file, with positions drawn as a Gaussian-mixture blob inside the sim-space box world2grid maps
# scenes to ([0.5, 1.5] of the [0, 2] grid, utils/transform_utils.py:8-15).
CONFIG_SHAPES = {
    # configs/lego.json:9-50 (N from the LFS pointer size / 248 B)
    "lego-like": dict(n=240_549, n_grid=50, material="jelly", E=2e5, nu=0.2, density=200.0, gravity=(0.0, 0.0, -100.0),
                      substep_dt=1e-4, frame_dt=1e-2, bcs=[
                          {"id": 0, "type": "fixed_cube", "center": [1.0, 1.2, 0.5], "size": [1.0, 0.8, 0.3], "start_time": 0, "num_dt": 10 ** 12},
                          {"id": 3, "type": "fixed_cube", "center": [0.0, 1.0, 1.0], "size": [0.1, 2.0, 2.0], "start_time": 0, "num_dt": 1},
                          {"id": 2, "type": "impulse", "center": [1.0, 0.65, 1.22], "size": [1.4, 0.18, 0.27], "force": [-0.8, 0.0, 0.0],
                           "start_time": 0.8, "num_dt": 10}]),
    # configs/cake_fracture_vertical.json
    "cake-like": dict(n=669_340, n_grid=50, material="plasticine", E=7000.0, nu=0.1, density=100.0, gravity=(0.0, -320.0, 0.0),
                      substep_dt=1e-4, frame_dt=1e-3, bcs=[
                          {"id": 0, "type": "fixed_cube", "center": [1.0, 1.4, 1.0], "size": [1.0, 0.2, 1.0], "start_time": 0, "num_dt": 10 ** 13},
                          {"id": 1, "type": "fixed_cube", "center": [1.0, 0.0, 1.0], "size": [1.0, 0.2, 1.0], "start_time": 0, "num_dt": 10 ** 13}]),
    # configs/metal_cupnoodle.json (N unknown: the .ply is a missing blob; 1 M assumed)
    "cupnoodle-like": dict(n=1_000_000, n_grid=100, material="metal", E=2e5, nu=0.4, density=500.0, gravity=(0.0, 20.0, 0.0),
                           substep_dt=1e-4, frame_dt=1e-2, bcs=[
                               {"id": 0, "type": "fixed_cube", "center": [2.0, 1.0, 1.0], "size": [0.4, 2.0, 2.0], "start_time": 0, "num_dt": 1},
                               {"id": 1, "type": "impulse", "center": [1.0, 1.5, 1.0], "size": [0.8, 0.8, 0.8], "force": [0.0, 0.003, 0.0],
                                "start_time": 0, "num_dt": 10 ** 16},
                               {"id": 4, "type": "fixed_cube", "center": [0.0, 1.0, 1.0], "size": [0.4, 2.0, 2.0], "start_time": 0, "num_dt": 1},
                               {"id": 5, "type": "fixed_cube", "center": [1.0, 2.0, 1.0], "size": [2.0, 0.4, 2.0], "start_time": 0, "num_dt": 1},
                               {"id": 6, "type": "fixed_cube", "center": [1.0, 0.0, 1.0], "size": [2.0, 0.4, 2.0], "start_time": 0, "num_dt": 1},
                               {"id": 7, "type": "fixed_cube", "center": [1.0, 1.0, 2.0], "size": [2.0, 2.0, 0.4], "start_time": 0, "num_dt": 1},
                               {"id": 8, "type": "fixed_cube", "center": [1.0, 1.0, 0.0], "size": [2.0, 2.0, 0.4], "start_time": 0, "num_dt": 1}]),
    # configs/bicycle.json (upper bound of N: every Gaussian of the scene inside sim_area)
    "bicycle-like": dict(n=3_616_103, n_grid=50, material="jelly", E=2e5, nu=0.4, density=100.0, gravity=(0.0, 0.0, -3.0),
                         substep_dt=1e-4, frame_dt=1e-2, bcs=[]),
}


def make_blob(n, n_grid, grid_extent=2.0, seed=SEED, n_components=24):
    """n particles as a Gaussian mixture inside [0.55, 1.45]^3 (x,y) x [0.6, 1.45] (z) of the [0, 2] grid: dense
    cells, hundreds of particles per cell at n_grid = 50 as in the named scenes; at rest; small random covariances"""
    rng = np.random.default_rng(seed)
    centres = rng.uniform(0.7, 1.3, (n_components, 3))
    sig = rng.uniform(0.04, 0.12, (n_components, 1))
    comp = rng.integers(0, n_components, n)
    xyz = centres[comp] + sig[comp] * rng.standard_normal((n, 3))
    xyz[:, :2] = np.clip(xyz[:, :2], 0.55, 1.45)
    xyz[:, 2] = np.clip(xyz[:, 2], 0.6, 1.45)
    dx = grid_extent / n_grid
    s2 = (rng.uniform(0.02, 0.2, (n, 3)) * dx) ** 2
    cov = np.stack([s2[:, 0], np.zeros(n), np.zeros(n), s2[:, 1], np.zeros(n), s2[:, 2]], 1)
    f = lambda x: np.ascontiguousarray(x, dtype=np.float32)
    return {"xyz": f(xyz), "vel": f(np.zeros((n, 3))), "cov": f(cov)}


def config_workload(name, seed=SEED):
    """(cloud without volumes, sim_args, ops) of a config-shaped case; volumes come from get_particle_volume as in
    main.py:270.  The collider is the one main.py:276 always adds (plane z = 0.4, normal +z, friction 0)."""
    c = CONFIG_SHAPES[name]
    cloud = make_blob(c["n"], c["n_grid"], seed=seed)
    args = sim_args(n_grid=c["n_grid"], grid_extent=2.0, material=c["material"], E=c["E"], nu=c["nu"], density=c["density"],
                    gravity=c["gravity"], substep_dt=c["substep_dt"], frame_dt=c["frame_dt"])
    args.boundary_conditions = c["bcs"]
    ops = {"boundary_conditions": c["bcs"], "collider": {"point": (0.0, 0.0, 0.4), "normal": (0.0, 0.0, 1.0), "friction": 0.0}}
    return cloud, args, ops
