"""Per-stage device timings of one substep on a synthetic workload (kernel tuning aid; bench.py holds the
judged contract).  python -m harness.stagebench [--workload S-4M] [--material jelly] [--strict-q1]"""
import argparse
import ctypes as C
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
    ap.add_argument("--material", default="jelly")
    ap.add_argument("--strict-q1", action="store_true")
    ap.add_argument("--sort-full", action="store_true")
    ap.add_argument("--warm", type=int, default=30)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--tag", default=os.environ.get("MPMK_LIB", "default"))
    a = ap.parse_args()
    import torch
    from harness import synthetic
    from mpm_solver.solver import MPM_Simulator
    dev = torch.device("cuda", 0)
    t = lambda x: torch.from_numpy(x).to(dev)
    if a.workload in synthetic.CONFIG_SHAPES:
        from internel_filling.filling import get_particle_volume
        cloud, args, ops = synthetic.config_workload(a.workload)
        cloud["vol"] = get_particle_volume(t(cloud["xyz"]), args).cpu().numpy()   # main.py:270
    else:
        cloud, args, ops = synthetic.workload(a.workload)
        args.material = a.material
    sim = MPM_Simulator(t(cloud["xyz"]), t(cloud["cov"]), t(cloud["vol"]), args, strict_q1=a.strict_q1)
    if a.sort_full:
        sim.set_parameters(sort_full=True)
    sim.set_boundary_conditions(ops["boundary_conditions"], args)
    sim.add_surface_collider(ops["collider"]["point"], ops["collider"]["normal"], friction=ops["collider"]["friction"])
    sim.mpm_state.particle_vel.from_torch(t(cloud["vel"]))
    dt = args.substep_dt
    n = sim.n_particles
    sim.step(a.warm, dt)
    sim.step(a.reps, dt)   # captures the graph the timed call replays
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    sim.step(a.reps, dt)
    e1.record(stream)
    torch.cuda.synchronize()
    graph_ms = e0.elapsed_time(e1) / a.reps
    b = sim._backend
    trace = None
    if os.environ.get("MPMK_TRACE"):
        buf = (C.c_ulonglong * 16)()
        b.ctx.call("mpmk_debug_trace", C.cast(buf, C.c_void_p))
        t0 = buf[0]
        trace = {k: round((buf[i] - t0) / 1e3, 1) for k, i in (("p2g_done", 1), ("grid_done", 2), ("g2p_done", 3), ("side_clear", 5), ("side_flagscan", 8), ("side_radix", 9), ("side_movers", 10), ("side_sort", 6), ("side_lists", 7))}
    sops = sim._ops_at(sim.time)
    stages = ["sort", "prepare", "p2g", "grid", "g2p"]
    acc = {k: 0.0 for k in stages}
    sp = stream.cuda_stream
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(6)] for _ in range(a.reps)]
    for r in range(a.reps):
        e = evs[r]
        e[0].record(stream); b.ctx.call("mpmk_sort", sp)
        e[1].record(stream); b.ctx.call("mpmk_prepare_grid", sp)
        e[2].record(stream); b.ctx.call("mpmk_p2g", dt, C.byref(sops), sp)
        e[3].record(stream); b.ctx.call("mpmk_grid_update", dt, C.byref(sops), sp)
        e[4].record(stream); b.ctx.call("mpmk_g2p", dt, sp)
        e[5].record(stream)
    torch.cuda.synchronize()
    for r in range(a.reps):
        for i, k in enumerate(stages):
            acc[k] += evs[r][i].elapsed_time(evs[r][i + 1]) / a.reps
    out = {"tag": a.tag, "workload": a.workload, "material": args.material, "n": n, "graph_ms_per_substep": round(graph_ms, 4),
           "Gpss": round(n / graph_ms / 1e6, 3), "stage_ms": {k: round(v, 4) for k, v in acc.items()},
           "trace_us": trace, "counters": sim.counters(), "finite": bool(torch.isfinite(sim.export()[0]).all())}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
