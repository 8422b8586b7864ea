#!/usr/bin/env python
"""Headless counterpart of the reference's main.py `simulate()` (main.py:164-335): everything around the solver that
does not need the rasteriser or a camera -- config JSON -> parameter groups, Gaussians from a 3DGS PLY (or a synthetic
blob), pre-processing rotations, sim-area mask, world2grid, covariance scaling, particle volumes, boundary conditions,
the hard-coded z = 0.4 collider, the frame loop, export into the rasteriser's frame, optional per-frame PLY dumps.

    python simulate.py --config_path configs/lego.json [--ply point_cloud.ply | --synthetic 240549] [--frames 10]

`means3D, cov3D_precomp = solver.export_world(...)` are the tensors render_frame (main.py:108-157) would hand to
diff-gaussian-rasterization; rendering itself is out of scope here (no rasteriser in this image)."""
import json
import os
import sys
import time
from argparse import ArgumentParser

PKG = os.path.dirname(os.path.abspath(__file__))
if PKG not in sys.path:
    sys.path.insert(0, PKG)

import numpy as np
import torch

from arguments import from_config
from internel_filling.filling import get_particle_volume
from mpm_solver.solver import MPM_Simulator, ti
from utils import gaussian_utils, ply_io
from utils.transform_utils import (apply_cov_rotations, apply_rotations, generate_rotation_matrices, simulatable_mask, world2grid)


def load_gaussians(ply, synthetic_n, sim_area, device, seed=0):
    """(xyz [P,3], cov [P,6]) on the device: from a 3DGS PLY (main.py:32-47 load_model + get_covariance), or random
    Gaussians filling the sim area when no trained model is at hand"""
    if ply:
        g = ply_io.read_gaussian_ply(ply)
        xyz = torch.from_numpy(g["xyz"]).to(device)
        cov = gaussian_utils.covariance_from_scaling_rotation(torch.from_numpy(g["scaling"]).to(device), torch.from_numpy(g["rotation"]).to(device))
        return xyz, cov
    rng = np.random.default_rng(seed)
    lo, hi = np.asarray(sim_area[0], np.float32), np.asarray(sim_area[1], np.float32)
    centres = rng.uniform(lo * 0.6, hi * 0.6, size=(24, 3)).astype(np.float32)
    xyz = centres[rng.integers(0, 24, synthetic_n)] + (rng.standard_normal((synthetic_n, 3)) * 0.12 * (hi - lo)).astype(np.float32)
    xyz = np.clip(xyz, lo * 0.999, hi * 0.999).astype(np.float32)
    log_s = np.log(rng.uniform(0.002, 0.01, size=(synthetic_n, 3)) * float((hi - lo).max())).astype(np.float32)
    rot = rng.standard_normal((synthetic_n, 4)).astype(np.float32)
    t = lambda a: torch.from_numpy(a).to(device)
    return t(xyz), gaussian_utils.covariance_from_scaling_rotation(t(log_s), t(rot))


def simulate(model_args, sim_args, render_args, ply=None, synthetic_n=100_000, frames=None, quiet=False):
    ti.init(arch=ti.cuda)                                                                   # main.py:28
    device = torch.device("cuda", torch.cuda.current_device())
    xyz, cov = load_gaussians(ply, synthetic_n, sim_args.sim_area, device)
    rotation_matrices = generate_rotation_matrices([torch.tensor(0.0)], [0], device=device)  # main.py:175
    rotated = apply_rotations(xyz, rotation_matrices)                                       # :176
    mask = simulatable_mask(rotated, sim_args.sim_area)                                     # :189-194
    sim_means3D = rotated[mask]
    sim_covs = apply_cov_rotations(cov[mask], rotation_matrices)                            # :234-237
    grid_means, pos_center, scaling_modifier = world2grid(sim_means3D, sim_args)            # :241
    grid_covs = sim_covs * (scaling_modifier * scaling_modifier)                            # :242
    volumes = get_particle_volume(grid_means, sim_args)                                     # :270
    solver = MPM_Simulator(grid_means, grid_covs, volumes, sim_args)                        # :272
    solver.set_boundary_conditions(sim_args.boundary_conditions, sim_args)                  # :273
    solver.add_surface_collider((0.0, 0.0, 0.4), (0.0, 0.0, 1.0))                           # :276
    n_frames = frames if frames is not None else render_args.num_frames
    log = []
    for fid in range(1, n_frames + 1):                                                      # :304-322
        torch.cuda.synchronize(); t0 = time.time()
        solver.step(sim_args.steps_per_frame, sim_args.substep_dt)                          # :305-306 as one graph replay
        means3D, cov3D = solver.export_world(scaling_modifier, pos_center, rotation_matrices, mode="render")   # :308-313, :139-146
        torch.cuda.synchronize(); dt = time.time() - t0
        if render_args.save_pcd and ply and fid % render_args.save_pcd_interval == 0:      # :315-320
            world, _ = solver.export_world(scaling_modifier, pos_center, mode="grid2world")
            ply_io.write_deformed_ply(ply, os.path.join(render_args.output_path, "point_cloud", f"iteration_{fid}", "point_cloud.ply"),
                                      mask.cpu().numpy(), world.cpu().numpy())
        log.append(dt)
        if not quiet:
            print(f"frame {fid}: {sim_args.steps_per_frame} substeps of {solver.n_particles} particles in {dt * 1e3:.2f} ms "
                  f"({solver.n_particles * sim_args.steps_per_frame / dt / 1e9:.2f} G particle-substeps/s), t = {solver.time:.4f}", flush=True)
    return solver, means3D, cov3D, log


if __name__ == "__main__":
    cp = ArgumentParser(add_help=False)
    cp.add_argument("--config_path", type=str, required=True)
    cp.add_argument("--ply", type=str, default=None)
    cp.add_argument("--synthetic", type=int, default=100_000)
    cp.add_argument("--frames", type=int, default=None)
    a, rest = cp.parse_known_args()
    with open(a.config_path) as f:
        config = json.load(f)
    simulate(*from_config(config, rest), ply=a.ply, synthetic_n=a.synthetic, frames=a.frames)
