"""The caller side of the hot path (main.py:164-335 minus the rasteriser): parameter groups from a config JSON,
Gaussian covariances, and the headless frame loop."""
import json
import os
import types

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(os.path.dirname(HERE), "gaussian-splatting-mpm_b200")


def test_arguments_from_config():
    from arguments import from_config
    cfg = json.load(open(os.path.join(PKG, "configs", "lego_shape.json")))
    m, s, r = from_config(cfg)
    assert s.n_grid == 50 and s.material == "jelly" and s.gravity == [0.0, 0.0, -100.0] and s.density == 200.0
    assert s.steps_per_frame == int(s.frame_dt / s.substep_dt) == 100                       # arguments/__init__.py:83
    assert len(s.boundary_conditions) == 3 and s.boundary_conditions[2]["type"] == "impulse"
    assert s.viscosity == 0.05 and s.fitting is False and s.rotation_degree == [0.0, 0.0, 0.0]   # defaults the JSON leaves alone
    assert r.num_frames == 10 and r.save_pcd is False and m.loaded_iter == -1
    # command-line overrides, as main.py:352-353 parses the remaining argv
    _, s2, r2 = from_config(cfg, ["--n_grid", "64", "--num_frames", "3", "--save_pcd"])
    assert s2.n_grid == 64 and r2.num_frames == 3 and r2.save_pcd is True
    _, s3, _ = from_config({"mpm": {"frame_dt": 1e-3, "substep_dt": 1e-4}})
    assert s3.steps_per_frame == int(1e-3 / 1e-4) and s3.E == 2e6 and s3.nu == 0.4


def test_gaussian_covariance():
    from utils import gaussian_utils as G
    rng = np.random.default_rng(0)
    q = rng.standard_normal((200, 4)).astype(np.float32)
    ls = rng.uniform(-5, -3, (200, 3)).astype(np.float32)
    cov = G.covariance_from_scaling_rotation(torch.from_numpy(ls), torch.from_numpy(q)).numpy()
    R = G.build_rotation(torch.from_numpy(q)).numpy().astype(np.float64)
    assert np.allclose(np.einsum("nij,nkj->nik", R, R), np.eye(3), atol=1e-5) and np.allclose(np.linalg.det(R), 1.0, atol=1e-5)
    S = np.exp(ls.astype(np.float64))
    full = np.einsum("nij,nj,nkj->nik", R, S * S, R)
    want = full.reshape(-1, 9)[:, [0, 1, 2, 4, 5, 8]]
    assert np.allclose(cov, want, rtol=1e-4, atol=1e-12)
    # w = 1 quaternion is the identity; sigmoid opacity
    I = G.build_rotation(torch.tensor([[2.0, 0.0, 0.0, 0.0]])).numpy()
    assert np.array_equal(I[0], np.eye(3, dtype=np.float32))
    assert abs(G.opacity_activation(torch.tensor([0.0])).item() - 0.5) < 1e-7


@pytest.mark.gpu
def test_headless_frame_loop(cuda_dev, tmp_path):
    """simulate.py on a synthetic 3DGS PLY with the lego-shaped config: frames advance, the export lands in the world
    frame, the per-frame dump only moves the simulated Gaussians"""
    import simulate
    from arguments import from_config
    from utils import ply_io
    rng = np.random.default_rng(2)
    P = 30000
    f = lambda *s: rng.standard_normal(s).astype(np.float32)
    xyz = (f(P, 3) * 0.35).astype(np.float32)
    xyz[:500] += 5.0                                                   # outside sim_area: not simulated
    g = {"xyz": xyz, "features_dc": f(P, 1, 3), "features_rest": f(P, 15, 3), "opacity": f(P, 1),
         "scaling": np.log(rng.uniform(0.004, 0.02, (P, 3))).astype(np.float32), "rotation": f(P, 4)}
    ply = str(tmp_path / "model" / "point_cloud.ply")
    ply_io.write_gaussian_ply(ply, g)
    cfg = json.load(open(os.path.join(PKG, "configs", "lego_shape.json")))
    cfg["render"]["output_path"] = str(tmp_path / "out")
    m, s, r = from_config(cfg, ["--save_pcd", "--save_pcd_interval", "2"])
    solver, means3D, cov3D, log = simulate.simulate(m, s, r, ply=ply, frames=2, quiet=True)
    inside = np.logical_and((xyz <= 1.3).all(1), (xyz >= -1.3).all(1))
    assert solver.n_particles == int(inside.sum()) < P and abs(solver.time - 2 * 100 * 1e-4) < 1e-9
    assert means3D.shape == (solver.n_particles, 3) and cov3D.shape == (solver.n_particles, 6)
    assert torch.isfinite(means3D).all() and torch.isfinite(cov3D).all()
    dumped = ply_io.read_gaussian_ply(os.path.join(cfg["render"]["output_path"], "point_cloud", "iteration_2", "point_cloud.ply"))
    assert np.array_equal(dumped["xyz"][~inside], xyz[~inside]) and np.array_equal(dumped["rotation"], g["rotation"])
    moved = np.abs(dumped["xyz"][inside] - xyz[inside]).max(axis=1)
    assert 0.0 < np.median(moved) < 0.5                                # gravity -100 for 0.02 s: centimetres, not metres


def test_orbit_camera_matrices():
    """render.orbit_camera (main.py:84-106 modify_cam): the camera sits on the sphere at the stated azimuth / elevation,
    looks at the centre (which projects to the middle of the frame at depth = radius), keeps the frame's vertical axis
    up on screen, and has a right-handed (right, down, forward) basis"""
    import render as R
    centre = np.array([0.3, -0.2, 0.5])
    frame = np.eye(3)
    cam = R.orbit_camera(centre, frame, torch.device("cpu"), azimuth=130.0, elevation=10.0, radius=5.75, width=320, height=240, fov_x=0.8, fov_y=0.6)
    eye = cam.cam_center.numpy().astype(np.float64)
    d = eye - centre
    assert abs(np.linalg.norm(d) - 5.75) < 1e-5
    assert abs(np.degrees(np.arcsin(d[2] / 5.75)) - 10.0) < 1e-4 and abs(np.degrees(np.arctan2(d[1], d[0])) - 130.0) < 1e-4
    V = cam.view_mat.numpy().T.astype(np.float64)                      # world -> camera (stored transposed, as the rasteriser wants)
    Rw = V[:3, :3]
    assert np.allclose(Rw @ Rw.T, np.eye(3), atol=1e-6) and abs(np.linalg.det(Rw) - 1.0) < 1e-6
    pc = V @ np.append(centre, 1.0)
    assert np.allclose(pc[:2], 0.0, atol=1e-5) and abs(pc[2] - 5.75) < 1e-5          # on the optical axis, in front
    up = V @ np.append(centre + np.array([0.0, 0.0, 1.0]), 1.0)
    assert up[1] < -0.5 and abs(up[0]) < 1e-5                                          # world up = image up (y points down)
    clip = np.append(centre, 1.0) @ cam.full_proj_mat.numpy().astype(np.float64)      # row vector x (view^T proj^T)
    ndc = clip[:3] / clip[3]
    assert np.allclose(ndc[:2], 0.0, atol=1e-5) and 0.0 < ndc[2] < 1.0
    side = np.append(centre, 1.0) + np.append(Rw[0] * 5.75 * np.tan(0.4), 0.0)         # a point on the right edge of the frustum
    c2 = side @ cam.full_proj_mat.numpy().astype(np.float64)
    assert abs(c2[0] / c2[3] - 1.0) < 1e-4


@pytest.mark.gpu
def test_frame_loop_renders_with_unchanged_rasterizer(cuda_dev, tmp_path):
    """simulate(..., render=True): every frame goes export_world -> the reference's unchanged rasteriser -> PNG; the falling
    blob is in view and the picture changes from frame to frame"""
    import render as R
    try:
        R.load_rasterizer()
    except RuntimeError as e:
        pytest.skip(str(e))
    import simulate
    from arguments import from_config
    from PIL import Image
    cfg = json.load(open(os.path.join(PKG, "configs", "lego_shape.json")))
    m, s, r = from_config(cfg, ["--num_frames", "3", "--output_path", str(tmp_path)])
    solver, means3D, cov3D, log = simulate.simulate(m, s, r, synthetic_n=20000, quiet=True, render=True)
    b = solver._backend
    assert means3D.data_ptr() == b.world_xyz.data_ptr() and cov3D.data_ptr() == b.world_cov.data_ptr()     # zero-copy hand-off
    frames = [np.asarray(Image.open(os.path.join(str(tmp_path), "images", f"{i:04d}.png")), dtype=np.int32) for i in (1, 2, 3)]
    assert all(f.shape == (800, 800, 3) for f in frames)
    assert all((f.max(2) > 12).mean() > 0.01 for f in frames)                          # something is on screen
    assert np.abs(frames[2] - frames[0]).max() > 8                                     # and it moves
