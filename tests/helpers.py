"""Shared builders for the parity tests: the same seeded inputs go to the CUDA solver (through the
drop-in Python API / C ABI) and to the CPU oracle."""
import numpy as np

from harness import synthetic
from oracle import mpm_oracle as O


def small_cloud(n_grid=32, cells_lo=(10, 10, 10), cells_n=(8, 8, 8), ppc=8, seed=7, material="jelly", **kw):
    cloud = synthetic.make_cloud(n_grid, 2.0, cells_lo, cells_n, ppc=ppc, seed=seed)
    args = synthetic.sim_args(n_grid=n_grid, material=material, **kw)
    return cloud, args


def random_state(n, seed, f_scale=0.05, stress_scale=50.0, c_scale=2.0):
    rng = np.random.default_rng(seed)
    C = (c_scale * rng.standard_normal((n, 3, 3))).astype(np.float32)
    F = (np.eye(3)[None] + f_scale * rng.standard_normal((n, 3, 3))).astype(np.float32)
    s = stress_scale * rng.standard_normal((n, 3, 3))
    stress = (0.5 * (s + s.transpose(0, 2, 1))).astype(np.float32)
    return C, F, stress


def make_oracle(cloud, args, ops=None, vel=True, **kw):
    sim = O.OracleSim(cloud["xyz"], cloud["cov"], cloud["vol"], args, **kw)
    if vel:
        sim.particle_vel[:] = cloud["vel"]
    if ops:
        sim.set_boundary_conditions(ops["boundary_conditions"], args)
        c = ops["collider"]
        sim.add_surface_collider(c["point"], c["normal"], friction=c["friction"])
    return sim


def make_gpu(cloud, args, dev, ops=None, vel=True, **kw):
    import torch
    from mpm_solver.solver import MPM_Simulator
    t = lambda a: torch.from_numpy(a).to(dev)
    sim = MPM_Simulator(t(cloud["xyz"]), t(cloud["cov"]), t(cloud["vol"]), args, **kw)
    if vel:
        sim.mpm_state.particle_vel.from_torch(t(cloud["vel"]))
    if ops:
        sim.set_boundary_conditions(ops["boundary_conditions"], args)
        c = ops["collider"]
        sim.add_surface_collider(c["point"], c["normal"], friction=c["friction"])
    return sim


def rel_fro(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))
