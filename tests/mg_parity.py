"""Run under torchrun with N >= 2 ranks (one GPU each): the slab-decomposed simulation against the same
simulation on one GPU and against the CPU oracle, canonicalised by particle id.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 tests/mg_parity.py
"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-mpm_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

from harness import synthetic  # noqa: E402
from mpm_solver import slabs  # noqa: E402
from mpm_solver.solver import MPM_Simulator  # noqa: E402

TOL_X, TOL_F = 1e-4, 1e-4   # max|dx| / grid dx ; relative Frobenius error of cov (F Sigma0 F^T)


def run_case(material, S, rank, world, dev, omega, ghost=1, migrate_every=5, out=sys.stdout, recut_at=None, skew=False):
    """recut_at: substep after which the slabs are re-cut at the particle-count quantiles (SlabSimulator.recut);
    skew: start from cuts that leave rank 0 with most of the block, so that the re-cut has something to balance"""
    n_grid, lo, cn, dt = 64, (22, 22, 22), (20, 20, 20), 4e-4
    if world > 2:   # at least 8 planes per slab (a slab must be >= 2*ghost+2 planes thick)
        cn = (8 * world, 16, 16)
        if lo[0] + cn[0] + 4 > n_grid:
            n_grid, lo, dt = 128, (30, 50, 50), 2e-4
    cloud = synthetic.make_cloud(n_grid, 2.0, lo, cn, seed=5, omega=omega)
    args = synthetic.sim_args(n_grid=n_grid, material=material, substep_dt=dt)
    ops = synthetic.block_ops(n_grid, 2.0, lo, cn)
    N = cloud["xyz"].shape[0]
    bounds = slabs.uniform_bounds(n_grid, lo[0], lo[0] + cn[0], world)
    if skew:   # all interior cuts pushed towards the high end of the block, tile-aligned, slabs still >= 2*ghost+2 planes
        hi = ((lo[0] + cn[0]) // 4) * 4
        bounds = [0] + [hi - 8 * (world - 1 - r) for r in range(1, world)] + [n_grid]
    part, ids = slabs.partition_cloud(cloud, bounds, rank, args.n_grid / args.grid_extent, n_grid)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    sl = slabs.SlabSimulator(t(part["xyz"]), t(part["cov"]), t(part["vol"]), args, t(ids), N, bounds, vel=t(part["vel"]),
                             ghost=ghost, migrate_every=migrate_every, capacity=(N if skew else None), boundary_conditions=ops["boundary_conditions"],
                             collider=ops["collider"], strict_q1=False)
    assert sl.global_count() == N
    balance = None
    if recut_at is None:
        sl.step(S, dt)
    else:
        sl.step(recut_at, dt)
        before = sl.imbalance()
        bounds = sl.recut()
        balance = (before, sl.imbalance())
        assert sl.global_count() == N, "particles were lost or duplicated in the re-cut"
        sl.step(S - recut_at, dt)
    assert sl.global_count() == N, "particles were lost or duplicated in migration"
    # ownership: every particle sits on the rank owning its base-cell plane right after a migration
    if sl._since_migration:
        sl.migrate()
    gid, _, _, _ = sl.gather()
    bx, _, _ = slabs.key_base_xyz(sl.backend.keys[: sl.n_particles], sl.nt)
    assert bool(((bx >= bounds[rank]) & (bx < bounds[rank + 1])).all())
    all_ids = [torch.zeros(0, dtype=torch.int64, device=dev)]
    cnt = torch.tensor([gid.numel()], device=dev)
    cnts = [torch.zeros_like(cnt) for _ in range(world)]
    dist.all_gather(cnts, cnt)
    x_mg, cov_mg, _ = (v.clone() for v in sl.export())
    moved = torch.tensor([sl.stats["sent"]], device=dev)
    dist.all_reduce(moved)
    ok = True
    msg = ""
    res = {"ranks": world, "particles": N, "substeps": S, "material": material, "migrated": int(moved.item())}
    if rank == 0:
        one = MPM_Simulator(t(cloud["xyz"]), t(cloud["cov"]), t(cloud["vol"]), args, strict_q1=False)
        one.mpm_state.particle_vel.from_torch(t(cloud["vel"]))
        one.set_boundary_conditions(ops["boundary_conditions"], args)
        one.add_surface_collider(ops["collider"]["point"], ops["collider"]["normal"], friction=ops["collider"]["friction"])
        one.step(S, dt)
        x1, c1, _ = one.export()
        dx = args.grid_extent / n_grid
        ex = float((x_mg - x1).abs().max() / dx)
        ec = float(torch.linalg.norm(cov_mg - c1) / torch.linalg.norm(c1))
        from oracle import mpm_oracle as O
        ref = O.OracleSim(cloud["xyz"], cloud["cov"], cloud["vol"], args, strict_q1=False)
        ref.particle_vel[:] = cloud["vel"]
        ref.set_boundary_conditions(ops["boundary_conditions"], args)
        ref.add_surface_collider(ops["collider"]["point"], ops["collider"]["normal"], friction=ops["collider"]["friction"])
        for _ in range(S):
            ref.p2g2p(dt)
        ref.postprocess()
        xo = ref.in_original_order("particle_xyz")
        co = ref.in_original_order("particle_cov").reshape(N, 6)
        exo = float(np.abs(x_mg.cpu().numpy() - xo).max() / dx)
        eco = float(np.linalg.norm(cov_mg.cpu().numpy() - co) / np.linalg.norm(co))
        msg = (f"{material}: N={N} ranks={world} S={S} migrated={int(moved.item())} counts={[int(c.item()) for c in cnts]} "
               f"vs 1 GPU: max|dx|/dx={ex:.2e} rel cov={ec:.2e}; vs oracle: {exo:.2e} {eco:.2e}")
        ok = ex < TOL_X and ec < TOL_F and exo < TOL_X and eco < TOL_F and int(moved.item()) > 0
        res.update({"max_dx_over_dx": max(ex, exo), "rel_cov": max(ec, eco), "vs_one_gpu": [ex, ec], "vs_oracle": [exo, eco],
                    "halo_path": sl.halo_path})
        if balance is not None:
            msg += f"; re-cut after {recut_at} substeps: max/mean count {balance[0]:.2f} -> {balance[1]:.2f}, bounds {bounds}"
            # (tile-aligned cuts at least 8 planes wide on a block of 20-64 planes: 1.05 at 2 ranks, 1.31 at 4)
            ok = ok and balance[1] < balance[0] and balance[1] < 1.5
            res["recut"] = {"at": recut_at, "imbalance_before": balance[0], "imbalance_after": balance[1]}
        print(msg, file=out, flush=True)
    flag = torch.tensor([1 if ok else 0], device=dev)
    dist.broadcast(flag, src=0)
    sl.close()
    res["ok"] = bool(flag.item())
    return res


def main():
    rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    ok = True
    only_recut = os.environ.get("MG_PARITY_ONLY") == "recut"
    for material, S, omega, ghost, every in () if only_recut else (("jelly", 40, (0.0, 0.0, 20.0), 1, 5), ("sand", 20, (0.0, 0.0, 20.0), 1, 5),
                                             ("metal", 30, (0.0, 0.0, 20.0), 2, "auto")):
        ok = run_case(material, S, rank, world, dev, omega, ghost, every)["ok"] and ok
    # a skewed start re-cut mid-run: same trajectory as one GPU and the oracle, counts balanced afterwards
    ok = run_case("metal", 30, rank, world, dev, (0.0, 0.0, 20.0), 2, "auto", recut_at=12, skew=True)["ok"] and ok
    dist.barrier()
    if rank == 0:
        print("mg parity ok" if ok else "mg parity FAILED", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
