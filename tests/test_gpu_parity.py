"""T1/T2: CUDA kernels (through the drop-in API and the C ABI) against the CPU oracle on the same
seeded inputs.  Integer work (cell ids, keys, sort permutation) is bit-exact; floating point is
compared with the tolerances written in each test."""
import ctypes as C

import numpy as np
import pytest

from harness import synthetic
from oracle import mpm_oracle as O
from helpers import small_cloud, make_oracle, make_gpu, random_state, rel_fro

pytestmark = pytest.mark.gpu


def _stream():
    import torch
    return torch.cuda.current_stream().cuda_stream


def _sorted_state(sim):
    """run sort + prepare on the GPU sim, return (sorted_keys, perm) as numpy uint32"""
    import torch
    b = sim._backend
    b.ctx.call("mpmk_sort", _stream())
    n = sim.n_particles
    sk = torch.empty(n, dtype=torch.int32, device=b.device)
    pm = torch.empty(n, dtype=torch.int32, device=b.device)
    b.ctx.call("mpmk_get_sort_result", sk.data_ptr(), pm.data_ptr(), _stream())
    torch.cuda.synchronize()
    return sk.cpu().numpy().view(np.uint32), pm.cpu().numpy().view(np.uint32)


@pytest.mark.parametrize("n_grid,lo,cn,ppc", [
    (50, (13, 13, 13), (12, 12, 12), 8),      # 18-bit keys: 2 passes x 9 bits; n_grid not a multiple of 4
    (100, (30, 30, 30), (10, 20, 30), 5),     # 20-bit keys: 3 passes x 7 bits
    (256, (100, 100, 100), (16, 16, 16), 8),  # 24-bit keys: 3 passes x 8 bits
    (512, (250, 250, 250), (10, 10, 10), 3),  # 27-bit keys: 3 passes x 9 bits
    (16, (5, 5, 5), (1, 1, 1), 1),            # a single particle
    (16, (5, 5, 5), (3, 3, 3), 13),           # ragged: not a multiple of the warp / chunk size
])
def test_keys_and_sort_bit_exact(cuda_dev, n_grid, lo, cn, ppc):
    cloud = synthetic.make_cloud(n_grid, 2.0, lo, cn, ppc=ppc, seed=11)
    rng = np.random.default_rng(2)
    shuffle = rng.permutation(len(cloud["xyz"]))
    for k in cloud:
        cloud[k] = np.ascontiguousarray(cloud[k][shuffle])
    args = synthetic.sim_args(n_grid=n_grid)
    sim = make_gpu(cloud, args, cuda_dev)
    import torch
    torch.cuda.synchronize()
    keys_gpu = sim._backend.keys[: sim.n_particles].cpu().numpy().view(np.uint32)
    _, keys_cpu = O.cell_keys(cloud["xyz"], n_grid, np.float32(n_grid / 2.0))
    assert np.array_equal(keys_gpu, keys_cpu)
    sk, pm = _sorted_state(sim)
    ref = np.argsort(keys_cpu, kind="stable").astype(np.uint32)
    assert np.array_equal(pm, ref)
    assert np.array_equal(sk, keys_cpu[ref])
    sk2, pm2 = O.stable_sort(keys_cpu)
    assert np.array_equal(pm, pm2) and np.array_equal(sk, sk2)


def test_sort_large_random_keys_bit_exact(cuda_dev):
    """4M-scale: sortedness + exact permutation against numpy's stable argsort"""
    cloud = synthetic.make_cloud(256, 2.0, (96, 96, 96), (64, 64, 64), ppc=4, seed=5)   # ~1.05 M
    rng = np.random.default_rng(9)
    sh = rng.permutation(len(cloud["xyz"]))
    for k in cloud:
        cloud[k] = np.ascontiguousarray(cloud[k][sh])
    sim = make_gpu(cloud, synthetic.sim_args(n_grid=256), cuda_dev)
    sk, pm = _sorted_state(sim)
    _, keys_cpu = O.cell_keys(cloud["xyz"], 256, np.float32(128.0))
    ref = np.argsort(keys_cpu, kind="stable").astype(np.uint32)
    assert np.array_equal(pm, ref)
    assert np.all(np.diff(sk.astype(np.int64)) >= 0)


def _set_state(gsim, osim, C_, F_, stress_, dev):
    import torch
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    gsim.mpm_state.particle_C.from_torch(t(C_)); osim.particle_C[:] = C_
    gsim.mpm_state.particle_F.from_torch(t(F_)); osim.particle_F[:] = F_; osim.particle_F_trial[:] = F_
    gsim.mpm_state.particle_stress.from_torch(t(stress_)); osim.particle_stress[:] = stress_


def test_p2g_grid_g2p_stagewise(cuda_dev):
    """each stage against the oracle's stage on identical inputs"""
    import torch
    cloud, args = small_cloud(n_grid=48, cells_lo=(14, 15, 16), cells_n=(10, 9, 11), ppc=6, material="jelly")
    ops = synthetic.block_ops(48, 2.0, (14, 15, 16), (10, 9, 11))
    g = make_gpu(cloud, args, cuda_dev, ops, strict_q1=False)
    o = make_oracle(cloud, args, ops, strict_q1=False)
    n = g.n_particles
    C_, F_, S_ = random_state(n, 21)
    _set_state(g, o, C_, F_, S_, cuda_dev)
    dt = 1e-4
    b = g._backend
    sops = g._ops_at(0.0)
    b.ctx.call("mpmk_sort", _stream())
    b.ctx.call("mpmk_prepare_grid", _stream())
    b.ctx.call("mpmk_p2g", dt, C.byref(sops), _stream())
    torch.cuda.synchronize()
    o._sort_and_reorder()
    o.p2g(dt)
    gm = g.mpm_state.grid_mass.to_torch().cpu().numpy()
    gv = g.mpm_state.grid_v_in.to_torch().cpu().numpy()
    # grid mass / momentum: 1e-5 relative (summation order differs)
    assert rel_fro(gm, o.grid_mass) < 1e-5
    assert rel_fro(gv, o.grid_v_in) < 1e-5
    assert np.array_equal(gm > 0, o.grid_mass > 0)   # same set of touched nodes
    # grid pass on the GPU's own p2g result vs oracle on ITS p2g result
    b.ctx.call("mpmk_grid_update", dt, C.byref(sops), _stream())
    torch.cuda.synchronize()
    o.grid_ops(dt)
    go = g.mpm_state.grid_v_out.to_torch().cpu().numpy()
    assert rel_fro(go, o.grid_v_out) < 1e-5
    # feed the oracle the GPU's grid so g2p is compared on identical grids: 1e-6
    o.grid_v_out[:] = go
    b.ctx.call("mpmk_g2p", dt, _stream())
    torch.cuda.synchronize()
    o.g2p(dt)
    st = g.mpm_state
    for name, tol in (("particle_vel", 1e-6), ("particle_C", 2e-6), ("particle_F_trial", 1e-6)):
        a = getattr(st, name).to_torch().cpu().numpy()
        assert rel_fro(a, o.in_original_order(name)) < tol, name
    x = st._xyz_packed.to_torch().cpu().numpy()
    # x += dt v: one fp32 ulp of x (x ~ 1) is 6e-8; FMA contraction on the GPU may differ from the oracle by that ulp
    assert np.abs(x - o.in_original_order("particle_xyz")).max() <= 1.2e-7


@pytest.mark.parametrize("material", ["jelly", "metal", "sand", "foam", "fluid", "plasticine"])
@pytest.mark.parametrize("strict", [True, False])
def test_constitutive_against_oracle(cuda_dev, material, strict):
    """return mapping + stress on random F_trial (large enough strains that yielding fires)"""
    import torch
    cloud, args = small_cloud(n_grid=32, cells_n=(6, 6, 6), material=material, E=2e5, nu=0.3)
    g = make_gpu(cloud, args, cuda_dev, strict_q1=strict, strict_q3=strict)
    o = make_oracle(cloud, args, strict_q1=strict, strict_q3=strict)
    n = g.n_particles
    rng = np.random.default_rng(33)
    F = (np.eye(3)[None] + 0.15 * rng.standard_normal((n, 3, 3))).astype(np.float32)
    F[: n // 4] = (np.eye(3)[None] + 1e-3 * rng.standard_normal((n // 4, 3, 3))).astype(np.float32)  # elastic regime
    ys = np.full(n, 5e3, np.float32)
    g.mpm_state.particle_F.from_torch(torch.from_numpy(F).to(cuda_dev))
    g.mpm_model.yield_stress.from_torch(torch.from_numpy(ys).to(cuda_dev))
    o.particle_F_trial[:] = F
    o.yield_stress[:] = ys
    dt = 1e-4
    g._backend.ctx.call("mpmk_constitutive", dt, _stream())
    torch.cuda.synchronize()
    o.compute_stress(dt)
    Fg = g.mpm_state.particle_F.to_torch().cpu().numpy()
    Sg = g.mpm_state.particle_stress.to_torch().cpu().numpy()
    ysg = g.mpm_model.yield_stress.to_torch().cpu().numpy()
    ok = np.isfinite(o.particle_stress).all(axis=(1, 2)) & np.isfinite(o.particle_F).all(axis=(1, 2))
    assert ok.mean() > 0.7
    # NaN pattern (sand with det F <= 0, :50-53 unclamped logs) must agree too
    assert np.array_equal(np.isfinite(Sg).all(axis=(1, 2)), np.isfinite(o.particle_stress).all(axis=(1, 2)))
    if material == "foam" and strict:
        # Hadamard quirk Q3 depends on the SVD's U, V individually: compare where sigma are well separated
        s = np.linalg.svd(F.astype(np.float64), compute_uv=False)
        ok &= (np.minimum(s[:, 0] - s[:, 1], s[:, 1] - s[:, 2]) > 0.05)
    assert rel_fro(Fg[ok], o.particle_F[ok]) < 2e-5
    scale = float(np.abs(o.particle_stress[ok]).max())
    assert np.abs(Sg[ok] - o.particle_stress[ok]).max() <= 2e-4 * scale
    assert np.allclose(ysg[ok], o.yield_stress[ok], rtol=1e-4, atol=1e-3)
    if material in ("metal", "plasticine", "foam", "fluid", "sand"):
        assert (np.abs(o.particle_F[ok] - F[ok]).reshape(ok.sum(), -1).max(1) > 1e-4).mean() > 0.2  # mapping fired


# x within 1e-4 dx; F and cov within 1e-4 relative (Frobenius), S substeps: 100 elastic, 20 plastic
TRAJ = [("jelly", True, 100), ("jelly", False, 100), ("metal", True, 20), ("sand", True, 20), ("foam", True, 20),
        ("foam", False, 20), ("fluid", True, 20), ("plasticine", True, 20)]


@pytest.mark.parametrize("material,strict,S", TRAJ)
def test_trajectory_parity(cuda_dev, material, strict, S):
    cloud, args = small_cloud(n_grid=40, cells_lo=(12, 12, 12), cells_n=(10, 10, 10), material=material, E=2e5, nu=0.3,
                              gravity=(0.0, 0.0, -9.8))
    ops = synthetic.block_ops(40, 2.0, (12, 12, 12), (10, 10, 10))
    if material == "metal":
        ops["boundary_conditions"].append({"id": 5, "type": "impulse", "center": [1.0, 1.0, 1.0], "size": [0.2, 0.2, 0.2],
                                           "force": [0.0, 0.05, 0.0], "start_time": 3e-4, "num_dt": 6})
    g = make_gpu(cloud, args, cuda_dev, ops, strict_q1=strict, strict_q3=strict)
    o = make_oracle(cloud, args, ops, strict_q1=strict, strict_q3=strict)
    dt = 1e-4
    for s in range(S):
        g.p2g2p(dt)
        o.p2g2p(dt)
    g.postprocess(); o.postprocess()
    import torch
    torch.cuda.synchronize()
    st = g.mpm_state
    x = st.particle_xyz.to_torch().cpu().numpy()
    if material == "foam" and strict:
        # Quirk Q3 (Hadamard F) feeds U_ii V_ii of an SVD whose singular values are nearly equal at the
        # start (F ~ I): U, V are ill-conditioned there, so the literal reference behaviour is chaotic at
        # rounding level.  Trajectories can only be checked for sanity; the kernel itself is compared on
        # well-separated singular values in test_constitutive_against_oracle.
        assert np.isfinite(x).all() and np.abs(x - o.in_original_order("particle_xyz")).max() < 0.5 * float(o.dx)
        return
    assert np.abs(x - o.in_original_order("particle_xyz")).max() < 1e-4 * float(o.dx)
    Ft = st.particle_F_trial.to_torch().cpu().numpy()
    assert rel_fro(Ft, o.in_original_order("particle_F_trial")) < 1e-4
    cov = st.particle_cov.to_torch().cpu().numpy()
    assert rel_fro(cov, o.in_original_order("particle_cov")) < 1e-4
    R = st.particle_R.to_torch().cpu().numpy()
    assert np.abs(R - o.in_original_order("particle_R")).max() < 1e-3
    v = st.particle_vel.to_torch().cpu().numpy()
    assert rel_fro(v, o.in_original_order("particle_vel")) < 1e-4
    # ordering contract: same particle ids in the same physical slots after the same substeps
    ids = g._backend.ids().cpu().numpy()
    o._sort_and_reorder(); g._backend.ctx.call("mpmk_sort", _stream())
    assert np.array_equal(ids[_sorted_state(g)[1]], o.order)


def _sort_info(sim):
    coh, mov = C.c_int(0), C.c_int64(0)
    sim._backend.ctx.call("mpmk_sort_info", C.byref(coh), C.byref(mov), _stream())
    return coh.value, mov.value


@pytest.mark.parametrize("n_grid,lo,cn,ppc,omega,dt,material", [
    (64, (22, 22, 22), (20, 20, 20), 8, (0.0, 0.0, 2.0), 4e-4, "jelly"),     # slow spin: ~1 % of the keys change per substep
    (64, (22, 22, 22), (20, 20, 20), 8, (3.0, -4.0, 40.0), 4e-4, "sand"),    # fast: a large fraction moves, tiles are crossed
    (50, (13, 13, 13), (12, 12, 12), 40, (0.0, 9.0, 25.0), 2e-4, "jelly"),   # 40 per cell: long equal-key runs, n_grid % 4 != 0
    (16, (5, 5, 5), (3, 3, 3), 13, (0.0, 0.0, 30.0), 1e-3, "jelly"),         # ragged, less than one chunk
    (128, (44, 44, 44), (40, 40, 40), 8, (1.0, 2.0, 12.0), 2e-4, "jelly"),   # 512 k particles: 125 chained chunks
])
def test_coherent_sort_bit_exact(cuda_dev, n_grid, lo, cn, ppc, omega, dt, material):
    """After a g2p the library merges the few particles whose cell changed into the order the rest already has
    (csrc/sort.cu, coherent sort).  The permutation and the sorted keys must be those of a CPU stable sort of the
    same keys, bit for bit, substep after substep."""
    import torch
    cloud = synthetic.make_cloud(n_grid, 2.0, lo, cn, ppc=ppc, seed=21, omega=omega)
    args = synthetic.sim_args(n_grid=n_grid, material=material, substep_dt=dt)
    g = make_gpu(cloud, args, cuda_dev)
    g.set_parameters(sort_merge=True)     # the library only merges from 1.5 M particles on by itself
    n = g.n_particles
    b = g._backend
    seen_coherent, movers = 0, []
    for s in range(12):
        g.p2g2p(dt) if s % 3 else g.step(2, dt)     # eager substeps and graph batches both end in a g2p
        torch.cuda.synchronize()
        keys = b.keys[:n].cpu().numpy().view(np.uint32).copy()
        sk, pm = _sorted_state(g)
        coh, mov = _sort_info(g)
        seen_coherent += coh
        movers.append(mov)
        ref = np.argsort(keys, kind="stable").astype(np.uint32)
        assert np.array_equal(pm, ref), f"substep {s}: permutation differs (coherent={coh}, movers={mov})"
        assert np.array_equal(sk, keys[ref])
    assert seen_coherent == 12, "the merge path did not run"
    assert max(movers) > 0 and max(movers) < n


def test_coherent_sort_equals_full_sort_trajectory(cuda_dev):
    """same particle order => same trajectories: MPMK_SORT_FULL against the default (only the order of the float
    atomics between warps may differ)"""
    cloud, args = small_cloud(n_grid=40, cells_lo=(12, 12, 12), cells_n=(8, 8, 8), material="metal")
    ops = synthetic.block_ops(40, 2.0, (12, 12, 12), (8, 8, 8))
    a = make_gpu(cloud, args, cuda_dev, ops)
    b = make_gpu(cloud, args, cuda_dev, ops)
    a.set_parameters(sort_merge=True)
    b.set_parameters(sort_full=True)
    a.step(30, 1e-4); b.step(30, 1e-4)
    assert _sort_info(a)[0] == 1 and _sort_info(b)[0] == 0
    assert np.array_equal(a._backend.ids().cpu().numpy(), b._backend.ids().cpu().numpy())   # same particle in every slot
    xa, ca, _ = a.export(); xb, cb, _ = b.export()
    assert (xa - xb).abs().max().item() <= 1e-6   # a few ulp of x ~ 1 (observed: <= 2 ulp; the atomics' order varies run to run)
    assert rel_fro(ca.cpu().numpy(), cb.cpu().numpy()) < 5e-5


def test_substeps_graph_equals_loop(cuda_dev):
    cloud, args = small_cloud(n_grid=40, cells_lo=(12, 12, 12), cells_n=(8, 8, 8), material="metal")
    ops = synthetic.block_ops(40, 2.0, (12, 12, 12), (8, 8, 8))
    ops["boundary_conditions"].append({"id": 3, "type": "fixed_cube", "center": [1.0, 1.0, 0.7], "size": [0.1, 0.1, 0.1],
                                       "start_time": 5e-4, "num_dt": 7})
    a = make_gpu(cloud, args, cuda_dev, ops)
    b = make_gpu(cloud, args, cuda_dev, ops)
    for _ in range(25):
        a.p2g2p(1e-4)
    b.step(10, 1e-4); b.step(15, 1e-4)
    assert abs(a.time - b.time) < 1e-12
    xa, ca, _ = a.export(); xb, cb, _ = b.export()
    # same kernels, same order of launches; only the float atomics' order may differ
    assert (xa - xb).abs().max().item() <= 1e-6   # a few ulp of x ~ 1 (observed: <= 2 ulp)
    assert rel_fro(ca.cpu().numpy(), cb.cpu().numpy()) < 5e-5


def test_track_cov_matches_update_cov(cuda_dev):
    cloud, args = small_cloud(n_grid=32, cells_n=(6, 6, 6))
    g = make_gpu(cloud, args, cuda_dev, track_cov=True)
    o = make_oracle(cloud, args, track_cov=True)
    for _ in range(10):
        g.p2g2p(1e-4); o.p2g2p(1e-4)
    cov = g.mpm_state.particle_cov.to_torch().cpu().numpy()   # no postprocess(): running update_cov value
    assert rel_fro(cov, o.in_original_order("particle_cov")) < 1e-5


def test_full_size_properties(cuda_dev):
    """S-4M (BASELINE.json's single-GPU size): size-independent properties instead of the oracle"""
    import torch
    cloud, args, ops = synthetic.workload("S-4M")
    g = make_gpu(cloud, args, cuda_dev, ops, strict_q1=False)
    n = g.n_particles
    assert n == 4096000
    b = g._backend
    sops = g._ops_at(0.0)
    b.ctx.call("mpmk_sort", _stream())
    sk, pm = _sorted_state(g)
    assert np.all(np.diff(sk.astype(np.int64)) >= 0)                      # sortedness
    assert np.array_equal(np.sort(pm), np.arange(n, dtype=np.uint32))     # a permutation
    keys = b.keys[:n].cpu().numpy().view(np.uint32)
    same = sk[1:] == sk[:-1]
    assert np.all(pm[1:][same] > pm[:-1][same])                           # stability
    assert np.array_equal(keys[pm], sk)
    b.ctx.call("mpmk_prepare_grid", _stream())
    b.ctx.call("mpmk_p2g", args.substep_dt, C.byref(sops), _stream())
    torch.cuda.synchronize()
    gin = b.grid_in.double()
    mass_p = g.mpm_state.particle_mass.to_torch().double()
    assert abs(gin[..., 3].sum().item() - mass_p.sum().item()) < 1e-6 * mass_p.sum().item()   # mass conservation
    mom_p = (mass_p[:, None] * torch.from_numpy(cloud["vel"]).to(cuda_dev).double()).sum(0)
    mom_g = gin[..., :3].reshape(-1, 3).sum(0)
    scale = (mass_p[:, None] * torch.from_numpy(cloud["vel"]).to(cuda_dev).double().abs()).sum().item()
    assert (mom_g - mom_p).abs().max().item() < 1e-6 * scale                                   # momentum conservation
    b.ctx.call("mpmk_grid_update", args.substep_dt, C.byref(sops), _stream())
    b.ctx.call("mpmk_g2p", args.substep_dt, _stream())
    for _ in range(3):
        g.p2g2p(args.substep_dt)
    c = g.counters()
    assert c["n_clamped"] == 0 and c["n_tiles"] >= 8000
    x, cov, R = g.export()
    assert torch.isfinite(x).all() and torch.isfinite(cov).all()
    RtR = torch.einsum("nji,njk->nik", R, R)
    assert (RtR - torch.eye(3, device=cuda_dev)).abs().max().item() < 1e-4     # R in O(3)
    # the sort after a g2p takes the merge path at this size: bit-exact against numpy at the full 4.1 M keys
    keys = b.keys[:n].cpu().numpy().view(np.uint32).copy()
    sk, pm = _sorted_state(g)
    coh, movers = _sort_info(g)
    assert coh == 1 and 0 < movers < n // 4
    ref = np.argsort(keys, kind="stable").astype(np.uint32)
    assert np.array_equal(pm, ref) and np.array_equal(sk, keys[ref])


@pytest.mark.gpu
def test_lego_shaped_config_parity(cuda_dev):
    """configs/lego.json by shape: 240,549 particles packed into a few thousand cells of a 50^3 grid
    (hundreds per cell: the same-address contention regime of the named scenes), gravity -100, the config's
    boundary conditions (one of them active on substep 0 only) and main.py's collider; volumes from
    get_particle_volume as main.py:270 does.  20 substeps against the oracle."""
    import torch
    from internel_filling.filling import get_particle_volume
    from oracle import mpm_oracle as O
    cloud, args, ops = synthetic.config_workload("lego-like")
    vol = get_particle_volume(torch.from_numpy(cloud["xyz"]).to(cuda_dev), args)
    cloud["vol"] = vol.cpu().numpy()
    assert np.allclose(cloud["vol"], O.particle_volume(cloud["xyz"], args.n_grid, args.grid_extent), rtol=1e-6)
    g = make_gpu(cloud, args, cuda_dev, ops, strict_q1=False)
    o = make_oracle(cloud, args, ops, strict_q1=False)
    dt = args.substep_dt
    S = 20
    g.step(S, dt)
    for _ in range(S):
        o.p2g2p(dt)
    g.postprocess(); o.postprocess()
    st = g.mpm_state
    x = st.particle_xyz.to_torch().cpu().numpy()
    assert np.abs(x - o.in_original_order("particle_xyz")).max() < 1e-4 * float(o.dx)
    assert rel_fro(st.particle_F_trial.to_torch().cpu().numpy(), o.in_original_order("particle_F_trial")) < 1e-4
    assert rel_fro(st.particle_cov.to_torch().cpu().numpy(), o.in_original_order("particle_cov")) < 1e-4
    assert abs(g.time - o.time) < 1e-12


def test_read_small_matches_cpu_copy(cuda_dev):
    """mpmk_read_small (a kernel stores a few words into pinned host memory; include/mpmk.h): same values as .cpu() for the
    shapes the frame path reads back -- a bool flag, the int64 migration table, an odd number of int32 words -- and still right
    while a bulk download is in flight on another stream (the case it exists for)."""
    import torch
    cloud = synthetic.make_cloud(16, 2.0, (4, 4, 4), (4, 4, 4), seed=1)
    args = synthetic.sim_args(n_grid=16, material="jelly")
    g = make_gpu(cloud, args, cuda_dev)
    b = g._backend
    gen = torch.Generator(device="cpu").manual_seed(5)
    for t in (torch.tensor([True], device=cuda_dev), torch.randint(-2**40, 2**40, (8, 9), generator=gen).to(cuda_dev),
              torch.arange(7, dtype=torch.int32, device=cuda_dev), torch.rand(3, generator=gen).to(cuda_dev)):
        got = b.read_small(t)
        want = t.to(torch.int32).cpu() if t.dtype == torch.bool else t.cpu()
        assert got.dtype == want.dtype and got.shape == want.shape and torch.equal(got, want)
    big_d = torch.zeros(1 << 26, dtype=torch.float32, device=cuda_dev)
    big_h = torch.empty(1 << 26, dtype=torch.float32).pin_memory()
    side = torch.cuda.Stream(device=cuda_dev)
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        big_h.copy_(big_d, non_blocking=True)
    t = torch.arange(100, dtype=torch.int64, device=cuda_dev) * 3
    assert torch.equal(b.read_small(t), t.cpu())
    torch.cuda.synchronize()
