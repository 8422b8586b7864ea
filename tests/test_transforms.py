"""the coordinate changes either side of the solver (utils/transform_utils.py mirror) against the
numpy restatement in oracle/transforms.py, and the fused export_world kernel against the reference's per-frame chain
postprocess -> to_torch -> grid2world -> render_frame's undo transforms (main.py:308-313, :134-146)."""
import types

import numpy as np
import pytest
import torch

from oracle import transforms as OT
from utils import transform_utils as T


def _cloud(n=5000, seed=3):
    rng = np.random.default_rng(seed)
    p = (rng.standard_normal((n, 3)) * [0.4, 1.3, 0.7] + [2.0, -1.0, 0.5]).astype(np.float32)
    a = rng.standard_normal((n, 3, 3)).astype(np.float32) * 0.05
    cov = np.einsum("nij,nkj->nik", a, a)
    return p, OT.upper_from_mat(cov).astype(np.float32)


def test_transform_mirror_matches_oracle():
    p, cov = _cloud()
    args = types.SimpleNamespace(grid_extent=2.0)
    Rs_t = T.generate_rotation_matrices([30.0, -75.0, 10.0], [0, 2, 1], device="cpu")
    Rs = [OT.rotation_matrix(d, a) for d, a in zip([30.0, -75.0, 10.0], [0, 2, 1])]
    for a, b in zip(Rs_t, Rs):
        assert np.allclose(a.numpy(), b, atol=1e-7)
    tp = torch.from_numpy(p); tc = torch.from_numpy(cov)
    rp = T.apply_rotations(tp, Rs_t); rc = T.apply_cov_rotations(tc, Rs_t)
    assert np.allclose(rp.numpy(), OT.apply_rotations(p, Rs), rtol=1e-6, atol=1e-6)
    assert np.allclose(rc.numpy(), OT.apply_cov_rotations(cov, Rs), rtol=1e-5, atol=1e-8)
    g, c, s = T.world2grid(rp, args)
    og, oc, os_ = OT.world2grid(rp.numpy(), 2.0)
    assert np.array_equal(c.numpy(), oc) and np.float32(s.item()) == os_
    assert np.allclose(g.numpy(), og, rtol=0, atol=1e-6)
    assert g.min().item() >= 0.5 - 1e-6 and g.max().item() <= 1.5 + 1e-6        # half the grid per axis
    # round trips: grid2world undoes world2grid, the inverse rotations undo the rotations
    w, wc = T.grid2world(g, rc * (s * s), s, c, args)
    assert np.allclose(w.numpy(), rp.numpy(), atol=2e-6) and np.allclose(wc.numpy(), rc.numpy(), rtol=1e-5, atol=1e-9)
    ow, owc = OT.grid2world(og, rc.numpy() * (os_ * os_), os_, oc, 2.0)
    assert np.allclose(w.numpy(), ow, atol=1e-6) and np.allclose(wc.numpy(), owc, rtol=1e-6, atol=1e-10)
    assert np.allclose(T.apply_inverse_rotations(rp, Rs_t).numpy(), p, atol=5e-6)
    assert np.allclose(T.apply_inverse_cov_rotations(rc, Rs_t).numpy(), cov, rtol=1e-4, atol=1e-8)
    assert np.allclose(T.apply_inverse_rotations(rp, Rs_t).numpy(), OT.apply_inverse_rotations(rp.numpy(), Rs), atol=1e-6)
    # upper <-> full
    assert np.array_equal(T.get_upper_from_mat(T.get_mat_from_upper(tc)).numpy(), cov)
    # sim-area mask, main.py:189-194
    m = T.simulatable_mask(tp, [[1.5, -2.5, 0.0], [2.5, 0.5, 1.0]])
    want = np.logical_and((p <= np.array([2.5, 0.5, 1.0], np.float32)).all(1), (p >= np.array([1.5, -2.5, 0.0], np.float32)).all(1))
    assert np.array_equal(m.numpy(), want) and 0 < want.sum() < len(p)
    # composite inverse rotation and affine
    M = T.inverse_rotation_composite(Rs_t)
    assert np.allclose(p.astype(np.float64) @ np.eye(3), (rp.numpy().astype(np.float64) @ M), atol=5e-6)
    al, be, ga = T.world_affine("grid2world", s, c, 2.0)
    assert np.allclose(g.numpy().astype(np.float64) * al + be, w.numpy(), atol=2e-6) and np.isclose(ga, 1.0 / float(s) ** 2)


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["grid2world", "render"])
def test_export_world_matches_reference_chain(cuda_dev, mode):
    """deform a cloud for a few substeps, then compare the one-kernel export with the chain the reference runs per frame"""
    from harness import synthetic
    from helpers import make_gpu
    cloud = synthetic.make_cloud(40, 2.0, (12, 12, 12), (10, 10, 10), seed=9, omega=(2.0, -3.0, 25.0))
    args = synthetic.sim_args(n_grid=40, material="metal", substep_dt=2e-4)
    ops = synthetic.block_ops(40, 2.0, (12, 12, 12), (10, 10, 10))
    g = make_gpu(cloud, args, cuda_dev, ops)
    g.step(15, 2e-4)
    degs, axes = [20.0, -40.0], [2, 0]
    Rs_t = T.generate_rotation_matrices(degs, axes, device=cuda_dev)
    Rs = [OT.rotation_matrix(d, a) for d, a in zip(degs, axes)]
    pos_center = np.array([0.3, -1.2, 2.5], np.float32); s = np.float32(0.37)
    xw, cw = g.export_world(float(s), torch.from_numpy(pos_center).to(cuda_dev), Rs_t, mode=mode)
    xw = xw.cpu().numpy(); cw = cw.cpu().numpy()
    # the reference chain on the solver's own postprocess output
    g.postprocess()
    x = g.mpm_state.particle_xyz.to_torch().cpu().numpy(); cov = g.mpm_state.particle_cov.to_torch().cpu().numpy().reshape(-1, 6)
    mw, cvw = OT.grid2world(x, cov, s, pos_center, 2.0)
    if mode == "render":
        mw, cvw = OT.render_inputs(mw, cvw, Rs, pos_center, 1.0)
    assert np.abs(xw - mw).max() <= 1e-5 * max(1.0, np.abs(mw).max())
    assert np.linalg.norm(cw - cvw) <= 1e-5 * np.linalg.norm(cvw)
    # and the torch mirror gives the same
    tm, tcv = T.grid2world(torch.from_numpy(x), torch.from_numpy(cov), torch.tensor(s), torch.from_numpy(pos_center), args)
    assert np.allclose(tm.numpy(), OT.grid2world(x, cov, s, pos_center, 2.0)[0], atol=1e-6)
