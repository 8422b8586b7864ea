"""Per-frame export: the reference's chain -- postprocess(), Field.to_torch(), grid2world, render_frame's
undo transforms (main.py:308-313, :134-146) as separate torch launches -- against the one-kernel export_world().
python -m harness.exportbench [--workload S-4M]"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
PKG = os.path.join(ROOT, "gaussian-splatting-mpm_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="S-4M")
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    import torch
    from harness import synthetic
    from mpm_solver.solver import MPM_Simulator
    from utils import transform_utils as T
    dev = torch.device("cuda", 0)
    t = lambda x: torch.from_numpy(x).to(dev)
    if a.workload in synthetic.CONFIG_SHAPES:
        from internel_filling.filling import get_particle_volume
        cloud, args, ops = synthetic.config_workload(a.workload)
        cloud["vol"] = get_particle_volume(t(cloud["xyz"]), args).cpu().numpy()   # main.py:270
    else:
        cloud, args, ops = synthetic.workload(a.workload)
    sim = MPM_Simulator(t(cloud["xyz"]), t(cloud["cov"]), t(cloud["vol"]), args, strict_q1=False)
    sim.mpm_state.particle_vel.from_torch(t(cloud["vel"]))
    sim.step(10, args.substep_dt)
    Rs = T.generate_rotation_matrices([20.0, -40.0], [2, 0], device=dev)
    c = torch.tensor([0.3, -1.2, 2.5], device=dev)
    s = torch.tensor(0.37, device=dev)
    n = sim.n_particles

    def chain():
        sim.postprocess()
        m, cv = T.grid2world(sim.mpm_state.particle_xyz.to_torch(), sim.mpm_state.particle_cov.to_torch(), s, c, args)
        m = T.apply_inverse_rotations(T.undotransform2origin(T.undoshift2center111(m), 1.0, c), Rs)
        return m, T.apply_inverse_cov_rotations(cv / (1.0 * 1.0), Rs)

    def fused():
        return sim.export_world(0.37, c, Rs, mode="render")

    out = {}
    for name, fn in (("reference_chain_torch", chain), ("export_world", fused)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        out[name + "_ms"] = e0.elapsed_time(e1) / a.reps
    m0, c0 = chain()
    m1, c1 = fused()
    out["max_abs_dx"] = float((m0 - m1).abs().max())
    out["rel_cov"] = float((c0 - c1).norm() / c0.norm())
    # algorithmic bytes of the fused kernel: read x 16 + F 48 + meta 16 + init_cov 24, write 12 + 24 per particle
    out["export_world_GBs"] = n * 140 / (out["export_world_ms"] * 1e-3) / 1e9
    print(json.dumps({"workload": a.workload, "n": n, **out}))


if __name__ == "__main__":
    main()
