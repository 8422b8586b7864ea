#!/usr/bin/env python
"""Headless counterpart of the reference's main.py `simulate()` (main.py:164-335): config JSON -> parameter groups,
Gaussians from a 3DGS PLY (or a synthetic blob), pre-processing rotations, sim-area mask, world2grid, covariance scaling,
particle volumes, boundary conditions, the hard-coded z = 0.4 collider, the frame loop, export into the rasteriser's
frame, optional per-frame PLY dumps and -- with --render -- the frames themselves.

    python simulate.py --config_path configs/lego.json [--ply point_cloud.ply | --synthetic 240549] [--frames 10] [--render]

`means3D, cov3D_precomp = solver.export_world(...)` are the tensors render_frame (main.py:108-157) hands to
diff-gaussian-rasterization.  --render does exactly that with the reference's unchanged rasteriser (render.py; built from
the reference tree by baseline/build_rasterizer.py) and writes output_path/images/%04d.png; the camera is main.py's orbit
camera (modify_cam) around the centre of the simulated Gaussians, since no cameras.json comes with a synthetic cloud."""
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
        # appearance as GaussianModel exposes it (gaussian_model.py: get_features = cat(dc, rest), get_opacity = sigmoid)
        shs = torch.from_numpy(np.concatenate([g["features_dc"], g["features_rest"]], axis=1)).to(device)
        load_gaussians.appearance = (shs, torch.sigmoid(torch.from_numpy(g["opacity"]).to(device)), 3)
        return xyz, cov
    rng = np.random.default_rng(seed)
    lo, hi = np.asarray(sim_area[0], np.float32), np.asarray(sim_area[1], np.float32)
    centres = rng.uniform(lo * 0.6, hi * 0.6, size=(24, 3)).astype(np.float32)
    xyz = centres[rng.integers(0, 24, synthetic_n)] + (rng.standard_normal((synthetic_n, 3)) * 0.12 * (hi - lo)).astype(np.float32)
    xyz = np.clip(xyz, lo * 0.999, hi * 0.999).astype(np.float32)
    log_s = np.log(rng.uniform(0.002, 0.01, size=(synthetic_n, 3)) * float((hi - lo).max())).astype(np.float32)
    rot = rng.standard_normal((synthetic_n, 4)).astype(np.float32)
    t = lambda a: torch.from_numpy(a).to(device)
    # degree-0 spherical harmonics: a colour from the position, so that a rendered blob shows its deformation
    rgb = 0.5 + 0.5 * np.sin(6.0 * (xyz - lo) / (hi - lo) + np.array([0.0, 2.0, 4.0], np.float32))
    load_gaussians.appearance = (t(((rgb - 0.5) / 0.28209479177387814).astype(np.float32)).reshape(-1, 1, 3),
                                 torch.full((synthetic_n, 1), 0.8, device=device), 0)
    return t(xyz), gaussian_utils.covariance_from_scaling_rotation(t(log_s), t(rot))


def simulate(model_args, sim_args, render_args, ply=None, synthetic_n=100_000, frames=None, quiet=False, render=False):
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
    if render:                                                                              # main.py:244-267, 296-300
        import render as R
        rasterizer = R.load_rasterizer()
        shs, opacity, sh_degree = load_gaussians.appearance
        shs, opacity = shs[mask].contiguous(), opacity[mask].contiguous()
        cam = None
        bg = torch.ones(3, device=device) if render_args.white_background else torch.zeros(3, device=device)   # main.py:222
        images = os.path.join(render_args.output_path or ".", "images")
    for fid in range(1, n_frames + 1):                                                      # :304-322
        torch.cuda.synchronize(); t0 = time.time()
        solver.step(sim_args.steps_per_frame, sim_args.substep_dt)                          # :305-306 as one graph replay
        means3D, cov3D = solver.export_world(scaling_modifier, pos_center, rotation_matrices, mode="render")   # :308-313, :139-146
        if render:                                                                          # :322 render_frame + save_frame
            if cam is None:
                # the orbit camera of main.py:84-106 around what is actually rendered: render_frame's undo chain runs with its
                # default scaling_modifier = 1.0 (main.py:108,139-146), so the frame it draws in is not the loader's world frame
                lo_r, hi_r = means3D.min(0).values, means3D.max(0).values
                cam = R.orbit_camera((0.5 * (lo_r + hi_r)).cpu().numpy(), np.eye(3), device, radius=1.3 * float((hi_r - lo_r).norm()),
                                     width=800, height=800)
            image, _ = R.render_frame(cam, means3D, cov3D, shs, opacity, bg, sh_degree=sh_degree, rasterizer=rasterizer)
        torch.cuda.synchronize(); dt = time.time() - t0
        if render:
            R.save_png(image, os.path.join(images, f"{fid:04d}.png"))
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
    cp.add_argument("--render", action="store_true", help="rasterise every frame with the reference's unchanged rasteriser (baseline/_ref)")
    a, rest = cp.parse_known_args()
    with open(a.config_path) as f:
        config = json.load(f)
    simulate(*from_config(config, rest), ply=a.ply, synthetic_n=a.synthetic, frames=a.frames, render=a.render)
