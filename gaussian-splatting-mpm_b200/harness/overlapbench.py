"""Does a latency-bound kernel chain (sort + work list) run for free beside the substep's big kernels?  (tuning aid)

Two independent simulators on one GPU: A replays whole substeps from a CUDA graph on one stream, B runs only
sort + prepare_grid, eagerly, on another.  Reported: A alone, B alone, both at once (device time from the first launch
to the last completion).  Perfect overlap: both ~= max(A, B); none: A + B.
    python -m harness.overlapbench [--workload S-4M] [--iters 100]"""
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
    ap.add_argument("--iters", type=int, default=100)
    ap.add_argument("--b-full-sort", action="store_true", help="B re-sorts in full (default: B advances with g2p-free merges is impossible, so B keeps its keys: zero movers)")
    a = ap.parse_args()
    import torch
    from harness import synthetic
    from mpm_solver.solver import MPM_Simulator
    dev = torch.device("cuda", 0)
    t = lambda x: torch.from_numpy(x).to(dev)
    cloud, args, ops = synthetic.workload(a.workload)

    def make():
        sim = MPM_Simulator(t(cloud["xyz"]), t(cloud["cov"]), t(cloud["vol"]), args, strict_q1=False)
        sim.set_boundary_conditions(ops["boundary_conditions"], args)
        sim.add_surface_collider(ops["collider"]["point"], ops["collider"]["normal"], friction=ops["collider"]["friction"])
        sim.mpm_state.particle_vel.from_torch(t(cloud["vel"]))
        return sim

    A, B = make(), make()
    B.set_parameters(sort_full=True)
    dt = args.substep_dt
    K = a.iters
    sA, sB = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    with torch.cuda.stream(sA):
        A.step(30, dt); A.step(K, dt); A.step(K, dt)      # both ping-pong parities of the K-substep graph are captured
    with torch.cuda.stream(sB):
        B.step(4, dt)
    torch.cuda.synchronize()

    def run_a():
        with torch.cuda.stream(sA):
            A.step(K, dt)

    def run_b():
        b = B._backend
        for _ in range(K):
            b.ctx.call("mpmk_sort", sB.cuda_stream)
            b.ctx.call("mpmk_prepare_grid", sB.cuda_stream)

    def timed(fns):
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True)
        ends = []
        cur = torch.cuda.current_stream()
        e0.record(cur)
        sA.wait_event(e0); sB.wait_event(e0)
        for f, s in fns:
            f()
            e = torch.cuda.Event(enable_timing=True)
            e.record(s)
            ends.append(e)
        torch.cuda.synchronize()
        return [e0.elapsed_time(e) / K for e in ends]

    out = {"workload": a.workload, "iters": K}
    out["A_alone_ms"] = timed([(run_a, sA)])[0]
    out["B_alone_ms"] = timed([(run_b, sB)])[0]
    both = timed([(run_a, sA), (run_b, sB)])
    out["both_A_ms"], out["both_B_ms"] = both
    both = timed([(run_b, sB), (run_a, sA)])
    out["both_Bfirst_B_ms"], out["both_Bfirst_A_ms"] = both
    print(json.dumps({k: (round(v, 4) if isinstance(v, float) else v) for k, v in out.items()}), flush=True)


if __name__ == "__main__":
    main()
