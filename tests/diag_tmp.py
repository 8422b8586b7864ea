import os, sys
import numpy as np
import torch
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (ROOT, os.path.join(ROOT, "gaussian-splatting-mpm_b200"), HERE):
    sys.path.insert(0, p)
from harness import synthetic
from helpers import make_gpu
dev = torch.device("cuda", 0)
cloud, args, ops = synthetic.workload("S-4M")
for no_lag in (False, True):
    g = make_gpu(cloud, args, dev, ops, strict_q1=False)
    if no_lag:
        g.set_parameters(no_lag=True)
    for k in (1, 1, 2, 5, 20):
        g.step(k, args.substep_dt)
        c = g.counters()
        gout = g._backend.grid_out
        nz = int((gout.abs().amax(-1) > 0).sum().item())
        gin = g._backend.grid_in
        print("no_lag", no_lag, "after", k, c, "nonzero grid_out nodes", nz, "nonzero grid_in", int((gin[..., 3] > 0).sum().item()), flush=True)
