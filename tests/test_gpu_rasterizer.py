"""The solver's output handed to the reference's UNCHANGED diff-gaussian-rasterization (main.py:108-157): built as-is
for sm_100a by baseline/build_rasterizer.py into baseline/_ref/ (git-ignored, travels to the GPU box), never copied
into the repository.  Checks the hand-off contract north_star names:
  * zero copy: the tensors the rasteriser receives ARE the export kernel's persistent output buffers (same data_ptr),
    in the [N,3] / [N,6] layout forward.cu:205-208 reads (cov3D_precomp + 6 * idx);
  * a frame rendered from export_world() equals the frame rendered from the reference's own per-frame chain
    (postprocess -> to_torch -> grid2world -> undo chain of render_frame) on the same state;
  * stream ordering: the rasteriser launches on the legacy default stream (forward.cu:389,428); with the
    solver on a side stream the consumer has to wait for it, and then gets the same frame."""
import math
import os
import sys

import numpy as np
import pytest

from harness import synthetic
from helpers import make_gpu

pytestmark = pytest.mark.gpu
REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref")


def _rasterizer():
    if not os.path.isdir(os.path.join(REF, "diff_gaussian_rasterization")):
        pytest.skip("baseline/_ref/diff_gaussian_rasterization not built (python baseline/build_rasterizer.py)")
    if REF not in sys.path:
        sys.path.insert(0, REF)
    import diff_gaussian_rasterization as dgr
    return dgr


def _camera(dev, target, dist, width=160, height=120, fov=0.9):
    """view / projection matrices as main.py:88-106 builds them (getWorld2View2 + getProjectionMatrix, transposed);
    the camera sits `dist` away from `target` and looks at it"""
    import torch
    target = np.asarray(target, np.float64)
    eye = target + dist * np.array([0.15, -1.0, 0.25]) / np.linalg.norm([0.15, -1.0, 0.25])
    fwd = (target - eye) / np.linalg.norm(target - eye)
    right = np.cross(fwd, [0.0, 0.0, 1.0]); right /= np.linalg.norm(right)
    down = np.cross(fwd, right)
    R = np.stack([right, down, fwd], 1)                    # camera-to-world rotation (columns = camera axes)
    W2C = np.eye(4); W2C[:3, :3] = R.T; W2C[:3, 3] = -R.T @ eye
    znear, zfar = 0.01, 100.0
    t = math.tan(fov / 2)
    P = np.zeros((4, 4))
    P[0, 0] = 1 / t; P[1, 1] = 1 / (t * height / width); P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear); P[2, 3] = -(zfar * znear) / (zfar - znear)
    view = torch.tensor(W2C, dtype=torch.float32, device=dev).t().contiguous()
    proj = torch.tensor(P, dtype=torch.float32, device=dev).t().contiguous()
    full = (view.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0).contiguous()
    fovy = 2 * math.atan(t * height / width)
    return dict(width=width, height=height, tanfovx=t, tanfovy=math.tan(fovy / 2), view=view, full=full,
                campos=torch.tensor(eye, dtype=torch.float32, device=dev))


def _render(dgr, cam, means3D, covs, shs, opac, dev):
    import torch
    st = dgr.GaussianRasterizationSettings(image_height=cam["height"], image_width=cam["width"], tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"],
                                           bg=torch.zeros(3, device=dev), scale_modifier=1.0, viewmatrix=cam["view"], projmatrix=cam["full"],
                                           sh_degree=0, campos=cam["campos"], prefiltered=False, debug=False)
    img, radii = dgr.GaussianRasterizer(raster_settings=st)(means3D=means3D, means2D=None, shs=shs, colors_precomp=None, opacities=opac,
                                                          scales=None, rotations=None, cov3D_precomp=covs)
    return img, radii


def test_export_world_feeds_unchanged_rasterizer(cuda_dev):
    import torch
    from utils import transform_utils as T
    dgr = _rasterizer()
    n_grid, lo, cn = 48, (16, 16, 16), (14, 14, 14)
    cloud = synthetic.make_cloud(n_grid, 2.0, lo, cn, ppc=8, seed=11, omega=(0.0, 0.0, 6.0))
    args = synthetic.sim_args(n_grid=n_grid, material="jelly", substep_dt=2e-4)
    ops = synthetic.block_ops(n_grid, 2.0, lo, cn)
    g = make_gpu(cloud, args, cuda_dev, ops, strict_q1=False)
    n = g.n_particles
    rng = np.random.default_rng(3)
    shs = torch.from_numpy(rng.uniform(-1, 1.5, (n, 1, 3)).astype(np.float32)).to(cuda_dev)      # degree-0 SH colours
    opac = torch.from_numpy(rng.uniform(0.2, 0.9, (n, 1)).astype(np.float32)).to(cuda_dev)
    # the loader's transform (main.py:203-262): a rotation about z, world2grid's scale and centre
    Rs = [torch.tensor([[math.cos(0.3), -math.sin(0.3), 0.0], [math.sin(0.3), math.cos(0.3), 0.0], [0.0, 0.0, 1.0]], dtype=torch.float32, device=cuda_dev)]
    scale, centre = 0.8, torch.tensor([0.1, -0.2, 0.05], dtype=torch.float32, device=cuda_dev)
    side = torch.cuda.Stream(device=cuda_dev)
    with torch.cuda.stream(side):
        g.step(30, args.substep_dt)
        xw, cw = g.export_world(scale, centre, Rs, mode="render")
    b = g._backend
    assert xw.data_ptr() == b.world_xyz.data_ptr() and cw.data_ptr() == b.world_cov.data_ptr()       # zero copy
    assert tuple(xw.shape) == (n, 3) and tuple(cw.shape) == (n, 6) and xw.is_contiguous() and cw.is_contiguous()
    # the rasteriser runs on the legacy default stream: the consumer orders it after the solver's stream
    torch.cuda.current_stream().wait_stream(side)
    lo_w, hi_w = xw.min(0).values.cpu().numpy(), xw.max(0).values.cpu().numpy()    # (reads on the current stream: ordered after the wait)
    cam = _camera(cuda_dev, 0.5 * (lo_w + hi_w), 2.2 * float(np.linalg.norm(hi_w - lo_w)))
    img, radii = _render(dgr, cam, xw, cw, shs, opac, cuda_dev)
    torch.cuda.synchronize()
    img = img.detach().cpu().numpy()
    assert img.shape == (3, cam["height"], cam["width"]) and np.isfinite(img).all()
    assert (radii > 0).float().mean().item() > 0.9                     # the block is in view
    assert (img.max(0) > 0.05).mean() > 0.02                           # and covers a visible part of the frame

    # the reference chain on the same state: postprocess -> to_torch -> grid2world -> render_frame's undo chain
    g.postprocess()
    pos = g.mpm_state.particle_xyz.to_torch()
    cov = g.mpm_state.particle_cov.to_torch().view(-1, 6)
    means_w, covs_w = T.grid2world(pos, cov, scale, centre, args)
    # (render_frame is called with its default scaling_modifier = 1.0, main.py:108,139-145)
    means_r = T.apply_inverse_rotations(T.undotransform2origin(T.undoshift2center111(means_w), 1.0, centre), Rs)
    covs_r = T.apply_inverse_cov_rotations(covs_w / (1.0 * 1.0), Rs)
    img2, _ = _render(dgr, cam, means_r.contiguous(), covs_r.contiguous(), shs, opac, cuda_dev)
    torch.cuda.synchronize()
    img2 = img2.detach().cpu().numpy()
    assert np.abs(img - img2).max() < 2e-3, float(np.abs(img - img2).max())     # 1e-5 relative on the inputs (composite vs step-by-step rounding)

    # after a full synchronisation the hand-off gives the very same frame (nothing was read early above)
    img3, _ = _render(dgr, cam, xw, cw, shs, opac, cuda_dev)
    torch.cuda.synchronize()
    assert np.array_equal(img3.detach().cpu().numpy(), img)
