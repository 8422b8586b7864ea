"""The differentiable ("fitting") path on the GPU -- MPM_state_opt + p2g2p_forward / p2g2p_backward / postprocess_forward /
postprocess_backward / learn (mpm_solver/fitting.py on csrc/opt.cu's forward + adjoint kernels), driven like
extra.py:199-236 -- against the float64 oracle oracle/fitting_ref.py (itself pinned on the CPU by tests/test_fitting_oracle.py:
end-to-end autograd and finite differences).  Tolerances: forward state 1e-5 relative (fp32 against fp64 over 30 substeps),
gradients 2e-4 relative Frobenius (30 substeps of fp32 adjoints with float atomics; observed 1e-6 .. 2e-5)."""
import numpy as np
import pytest

from harness import synthetic
from helpers import rel_fro
from oracle.fitting_ref import FittingRef

pytestmark = pytest.mark.gpu
S = 30


def _setup(cuda_dev, strict_grad, with_box=True):
    import torch
    from mpm_solver.solver import MPM_Simulator
    from mpm_solver.boundary_conditions import boundaryConditionTypeCallBacks
    n_grid, lo, cn = 16, (5, 5, 5), (4, 4, 4)
    cloud = synthetic.make_cloud(n_grid, 2.0, lo, cn, ppc=8, seed=3, omega=(0.0, 0.0, 3.0))
    dx = 2.0 / n_grid
    vel = (cloud["vel"] + np.array([0.3, 0.0, -1.0], np.float32)).astype(np.float32)
    args = synthetic.sim_args(n_grid=n_grid, material="jelly", E=2e4, nu=0.3, density=1000.0, gravity=(0.0, 0.0, -9.8), substep_dt=1e-3)
    args.fitting = True
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(cuda_dev)
    sim = MPM_Simulator(t(cloud["xyz"]), t(cloud["cov"]), t(cloud["vol"]), args, init_v=t(vel), strict_grad=strict_grad)
    box = ([1.0, 1.0, lo[2] * dx - 0.5 * dx], [1.0, 1.0, 1.5 * dx])
    if with_box:
        bc = boundaryConditionTypeCallBacks["fixed_cube"](sim.n_particles, {"id": 0, "type": "fixed_cube", "center": box[0], "size": box[1],
                                                                             "start_time": 0.0, "num_dt": 10 ** 9}, args)
        sim.grid_postprocess.append(bc)
    else:
        sim.set_bc_ground_only()       # StickyGroundBC: center [1, 0.6, 1], size [1, 0.1, 1] (boundary_conditions.py:88-91), far from the block
        box = ([1.0, 0.6, 1.0], [1.0, 0.1, 1.0])
    ref = FittingRef(cloud["xyz"], cloud["cov"], cloud["vol"], vel, n_grid, 2.0, E=2e4, nu=0.3, density=1000.0, gravity=(0.0, 0.0, -9.8),
                     box=box, n_steps=S)
    return sim, ref


@pytest.mark.parametrize("strict_grad", [False, True])
def test_forward_and_gradients_against_float64_oracle(cuda_dev, strict_grad):
    import torch
    sim, ref = _setup(cuda_dev, strict_grad)
    dt = 1e-3
    for s in range(S):
        sim.p2g2p_forward(dt, s)
    sim.postprocess_forward()
    with torch.no_grad():
        H = ref.forward(dt)
    st = sim.mpm_state
    assert rel_fro(st.particle_xyz.to_torch()[S].cpu().numpy(), H["x"][S].numpy()) < 1e-6
    assert rel_fro(st.particle_F.to_torch()[S].cpu().numpy(), H["F"][S].numpy()) < 1e-5
    assert rel_fro(st.particle_C.to_torch()[S].cpu().numpy(), H["C"][S].numpy()) < 1e-4
    assert rel_fro(st.particle_stress.to_torch()[S - 1].cpu().numpy(), H["stress"][S - 1].numpy()) < 1e-3     # stress ~ E x strain: cancellation
    assert rel_fro(st.particle_cov.to_torch().cpu().numpy().reshape(-1, 6), H["cov"].numpy()) < 1e-5
    assert float((H["F"][S] - torch.eye(3, dtype=torch.float64)).abs().max()) > 1e-3       # the block really deformed
    # loss gradients as extra.py:214-221 hands them over
    rng = np.random.default_rng(0)
    g_x = rng.standard_normal((ref.n, 3)).astype(np.float32); g_cov = (rng.standard_normal((ref.n, 6)) * 10.0).astype(np.float32)
    sim.clear_grads()
    st.set_grads(g_x, g_cov.reshape(-1))
    sim.postprocess_backward()
    for s in reversed(range(S)):
        sim.p2g2p_backward(dt, s)
    gl, gy, gx0, extra = ref.backward(dt, g_x, g_cov, strict=strict_grad)
    md = sim.mpm_model
    e = {"logE": rel_fro(md.logE.grad.to_torch().cpu().numpy(), gl.numpy()), "y": rel_fro(md.y.grad.to_torch().cpu().numpy(), gy.numpy()),
         "x0": rel_fro(st.particle_xyz.grad.to_torch()[0].cpu().numpy(), gx0.numpy()),
         "mu": rel_fro(md.mu.grad.to_torch().cpu().numpy(), extra["g_mu"].numpy()),
         "F0": rel_fro(st.particle_F.grad.to_torch()[0].cpu().numpy(), extra["gF0"].numpy())}
    print("strict_grad", strict_grad, e)
    assert all(v < 2e-4 for v in e.values()), e
    assert float(gl.abs().max()) > 0


def test_learn_clips_and_steps(cuda_dev):
    import torch
    sim, _ = _setup(cuda_dev, True, with_box=False)
    md = sim.mpm_model
    n = sim.n_particles
    g = torch.linspace(-3.0, 3.0, n, device=cuda_dev)
    md.logE.grad.from_torch(g); md.y.grad.from_torch(-g)
    le0, y0 = md.logE.to_torch().clone(), md.y.to_torch().clone()
    sim.learn()
    assert torch.allclose(md.logE.to_torch(), le0 - 0.8 * g.clamp(-1, 1))          # solver.py:95-105
    assert torch.allclose(md.y.to_torch(), y0 + 1.6 * g.clamp(-1, 1))
    # cycle_init (model.py:221-228) and the p2g2p(dt, s) spelling of extra.py:200
    for s in range(S):
        sim.p2g2p(1e-3, s)
    sim.mpm_state.cycle_init()
    st = sim.mpm_state
    assert torch.equal(st.particle_xyz.to_torch()[0], st.particle_xyz.to_torch()[S])
    assert torch.equal(st.particle_F.to_torch()[0], st.particle_F.to_torch()[S])


def test_fitting_requires_a_box_as_first_grid_operation(cuda_dev):
    sim, _ = _setup(cuda_dev, True, with_box=False)
    sim.grid_postprocess.clear()
    with pytest.raises(IndexError):       # the reference indexes grid_postprocess[0] unconditionally (solver.py:66)
        sim.p2g2p_forward(1e-3, 0)
