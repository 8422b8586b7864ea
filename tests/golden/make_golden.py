"""Regenerates tests/golden/mpm_golden.npz from the CPU oracle (oracle/mpm_oracle.c).

The reference ships no golden vectors for this path and Taichi cannot be imported here (DESIGN.md 2:
parity unpinned), so these are ORACLE-generated regression vectors: they pin the oracle against
drift and give the -m gpu tests a committed target; they are not outputs of the reference.

    python tests/golden/make_golden.py
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-mpm_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

from harness import synthetic  # noqa: E402
from oracle import mpm_oracle as O  # noqa: E402

N_GRID, LO, CN, SEED, DT = 16, (5, 5, 5), (4, 4, 4), 99, 1e-3
CASES = [("jelly", True, 8), ("jelly", False, 8), ("metal", True, 8), ("sand", True, 8), ("foam", True, 8),
         ("foam", False, 8), ("fluid", True, 8), ("plasticine", True, 8)]


def inputs():
    cloud = synthetic.make_cloud(N_GRID, 2.0, LO, CN, seed=SEED, omega=(0.0, 0.0, 6.0))
    ops = synthetic.block_ops(N_GRID, 2.0, LO, CN)
    return cloud, ops


def run_case(cloud, ops, material, strict, S):
    args = synthetic.sim_args(n_grid=N_GRID, material=material, substep_dt=DT, E=5e4)
    sim = O.OracleSim(cloud["xyz"], cloud["cov"], cloud["vol"], args, strict_q1=strict, strict_q3=strict)
    sim.particle_vel[:] = cloud["vel"]
    sim.set_boundary_conditions(ops["boundary_conditions"], args)
    c = ops["collider"]
    sim.add_surface_collider(c["point"], c["normal"], friction=c["friction"])
    for _ in range(S):
        sim.p2g2p(DT)
    sim.postprocess()
    return {"xyz": sim.in_original_order("particle_xyz"), "F_trial": sim.in_original_order("particle_F_trial"),
            "cov": sim.in_original_order("particle_cov")}


def main():
    cloud, ops = inputs()
    out = {}
    inv_dx = N_GRID / 2.0
    _, keys = O.cell_keys(cloud["xyz"], N_GRID, inv_dx)
    skeys, perm = O.stable_sort(keys)
    out["keys"], out["sorted_keys"], out["perm"] = keys, skeys, perm
    for material, strict, S in CASES:
        r = run_case(cloud, ops, material, strict, S)
        tag = f"{material}_{'strict' if strict else 'intended'}"
        for k, v in r.items():
            out[f"{tag}_{k}"] = np.asarray(v, np.float32)
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "mpm_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, {k: v.shape for k, v in out.items() if k.endswith("_F_trial")})


if __name__ == "__main__":
    main()
