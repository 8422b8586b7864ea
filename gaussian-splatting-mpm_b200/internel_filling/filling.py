"""Particle volume from per-cell occupancy (reference: internel_filling/filling.py:12-42), without
Taichi: two small CUDA kernels in lib/libmpmk.so (csrc/export.cu k_vol_count / k_vol)."""
from __future__ import annotations

import torch

from mpm_solver import _native


def get_particle_volume(pos: torch.Tensor, args, uniform: bool = False):
    if not pos.is_cuda:
        raise _native.MpmkError("get_particle_volume needs a CUDA tensor; there is no CPU fallback")
    num_pts = pos.shape[0]
    pos = pos.detach().to(torch.float32).contiguous()
    n = int(args.n_grid)
    grid_dx = args.grid_extent / args.n_grid
    ctx = _native.Context(pos.device.index or 0, n, args.grid_extent, 0)
    try:
        grid = torch.zeros(n * n * n, dtype=torch.int32, device=pos.device)
        vol = torch.empty(num_pts, dtype=torch.float32, device=pos.device)
        ctx.call("mpmk_particle_volume", num_pts, pos.data_ptr(), float(grid_dx), grid.data_ptr(), vol.data_ptr(),
                 torch.cuda.current_stream().cuda_stream)
        torch.cuda.current_stream().synchronize()
    finally:
        ctx.close()
    if uniform:  # same volume for all particles
        return torch.mean(vol).repeat(num_pts)
    return vol
