"""Lagged substeps (csrc/api.cu): inside a step() batch the sort of substep s runs on a side branch of the graph while
p2g / grid of substep s consume the work list of substep s-1; g2p scatters every particle to the slot that sort ranks it
at.  Checked here:
  * the side-branch sort is still THE stable sort of the current keys, bit for bit (permutation, sorted keys, ranks =
    inverse permutation), and g2p really moved every particle to its rank;
  * particles that left their work item's tile between the two sorts ("crossers") exist in the fast clouds and are
    handled (scatter straight to the global grid + block list, gather from the global grid);
  * trajectories equal the sequential chain's (set_parameters(no_lag=True)) to rounding, and the oracle's to 1e-4."""
import ctypes as C

import numpy as np
import pytest

from harness import synthetic
from helpers import small_cloud, make_oracle, make_gpu, rel_fro

pytestmark = pytest.mark.gpu


def _stream():
    import torch
    return torch.cuda.current_stream().cuda_stream


def _last_sort(sim):
    import torch
    b = sim._backend
    n = sim.n_particles
    sk, pm, rk = (torch.empty(n, dtype=torch.int32, device=b.device) for _ in range(3))
    b.ctx.call("mpmk_get_last_sort", sk.data_ptr(), pm.data_ptr(), rk.data_ptr(), _stream())
    torch.cuda.synchronize()
    return tuple(t.cpu().numpy().view(np.uint32) for t in (sk, pm, rk))


@pytest.mark.parametrize("n_grid,lo,cn,ppc,omega,dt,material,merge", [
    (64, (22, 22, 22), (20, 20, 20), 8, (0.0, 0.0, 2.0), 4e-4, "jelly", True),      # slow spin, merge path on the side branch
    (64, (22, 22, 22), (20, 20, 20), 8, (3.0, -4.0, 40.0), 4e-4, "sand", True),     # fast: tiles are crossed every substep
    (64, (22, 22, 22), (20, 20, 20), 8, (3.0, -4.0, 40.0), 4e-4, "metal", False),   # same, full radix path + rank inversion
    (50, (13, 13, 13), (12, 12, 12), 40, (0.0, 9.0, 25.0), 2e-4, "jelly", True),    # 40 per cell, n_grid % 4 != 0
    (16, (5, 5, 5), (3, 3, 3), 13, (0.0, 0.0, 30.0), 1e-3, "jelly", False),         # ragged, less than one chunk
    (128, (44, 44, 44), (40, 40, 40), 8, (1.0, 2.0, 12.0), 2e-4, "jelly", True),    # 512 k particles
])
def test_lagged_sort_bit_exact_and_scatter(cuda_dev, n_grid, lo, cn, ppc, omega, dt, material, merge):
    import torch
    cloud = synthetic.make_cloud(n_grid, 2.0, lo, cn, ppc=ppc, seed=21, omega=omega)
    args = synthetic.sim_args(n_grid=n_grid, material=material, substep_dt=dt)
    g = make_gpu(cloud, args, cuda_dev)
    if merge:
        g.set_parameters(sort_merge=True)     # the library only merges from 1.5 M particles on by itself
    n = g.n_particles
    b = g._backend
    g.step(3, dt)     # one sequential substep (nothing is sorted yet), two lagged ones
    crossers = 0
    for s in range(8):
        torch.cuda.synchronize()
        keys = b.keys[:n].cpu().numpy().view(np.uint32).copy()       # current keys, physical order
        ids = b.ids().cpu().numpy().copy()                           # particle in each physical slot
        g.step(1, dt) if s % 2 else g.step(3, dt)                    # ends with a lagged substep either way
        if s % 2 == 0:
            continue   # only the last sort of a batch is visible: check the single-substep batches
        sk, pm, rk = _last_sort(g)
        ref = np.argsort(keys, kind="stable").astype(np.uint32)
        assert np.array_equal(pm, ref), f"substep {s}: side-branch sort differs from the stable sort of the keys"
        assert np.array_equal(sk, keys[ref])
        inv = np.empty(n, np.uint32); inv[ref] = np.arange(n, dtype=np.uint32)
        assert np.array_equal(rk, inv)
        assert np.array_equal(b.ids().cpu().numpy(), ids[ref]), "g2p did not move every particle to its rank"
        crossers += g.counters()["n_crossers"]
    if max(abs(w) for w in omega) >= 40.0:
        assert crossers > 0, "no particle ever left its tile: the crosser path did not run"
    x, cov, _ = g.export()
    assert torch.isfinite(x).all() and torch.isfinite(cov).all()


@pytest.mark.parametrize("material,omega,S", [("jelly", (0.0, 0.0, 2.0), 40), ("metal", (3.0, -4.0, 40.0), 24), ("sand", (0.0, 9.0, 25.0), 24)])
def test_lagged_equals_sequential_and_oracle(cuda_dev, material, omega, S):
    n_grid, lo, cn = 48, (14, 14, 14), (12, 12, 12)
    cloud = synthetic.make_cloud(n_grid, 2.0, lo, cn, ppc=8, seed=5, omega=omega)
    args = synthetic.sim_args(n_grid=n_grid, material=material, substep_dt=2e-4)
    ops = synthetic.block_ops(n_grid, 2.0, lo, cn)
    dx = 2.0 / n_grid
    ops["boundary_conditions"].append({"id": 2, "type": "impulse", "center": [1.0, 1.0, 1.0], "size": [0.1, 0.3, 0.3], "force": [0.02, 0.0, 0.01],
                                       "start_time": 6e-4, "num_dt": 9})
    a = make_gpu(cloud, args, cuda_dev, ops, strict_q1=False)
    b = make_gpu(cloud, args, cuda_dev, ops, strict_q1=False)
    b.set_parameters(no_lag=True)
    o = make_oracle(cloud, args, ops, strict_q1=False)
    dt = args.substep_dt
    for k in (S // 2, S - S // 2):
        a.step(k, dt); b.step(k, dt)
    for _ in range(S):
        o.p2g2p(dt)
    xa, ca, _ = a.export(); xb, cb, _ = b.export()
    o.postprocess()
    xo = o.in_original_order("particle_xyz")
    ea = float(np.abs(xa.cpu().numpy() - xo).max() / dx); eb = float(np.abs(xb.cpu().numpy() - xo).max() / dx)
    eab = float((xa - xb).abs().max().item() / dx)
    print(material, "lagged vs oracle", ea, "sequential vs oracle", eb, "lagged vs sequential", eab, "crossers", a.counters()["n_crossers"])
    assert ea < 1e-4 and eb < 1e-4 and eab < 1e-4
    assert rel_fro(ca.cpu().numpy().reshape(-1), o.in_original_order("particle_cov")) < 1e-4
    assert rel_fro(ca.cpu().numpy(), cb.cpu().numpy()) < 1e-4
    # the grid the batch leaves behind: velocities of the last substep on the nodes it touched, nothing else
    ga = a.mpm_state.grid_v_out.to_torch().cpu().numpy(); gb = b.mpm_state.grid_v_out.to_torch().cpu().numpy()
    assert np.array_equal(np.abs(ga).max(-1) > 0, np.abs(gb).max(-1) > 0) or rel_fro(ga, gb) < 1e-4
    assert rel_fro(ga, gb) < 1e-3
