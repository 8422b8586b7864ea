"""Names the reference's mpm_solver/utils.py exports that host code relies on.

The kernels themselves (compute_stress_from_F_trial :14-54, p2g :90-134,
grid_normalization_and_gravity :178-183, g2p :219-282, compute_R_from_F :377-398,
compute_cov_from_F :402-433) live in csrc/*.cu; here are the material table (:5-10) and the two
init helpers that model.py calls (:350-373), re-hosted on torch CUDA tensors.
"""
from __future__ import annotations

import torch

from . import _native

# mpm_solver/utils.py:5-10; "fluid" and "plasticine" are additions of this build
# (SURVEY.md 9.5): north_star asks for them, the reference raises TypeError for both.
material_types = {
    "jelly": 0,
    "metal": 1,
    "sand": 2,
    "foam": 3,
    "fluid": 4,
    "plasticine": 5,
}


def _stream():
    return torch.cuda.current_stream().cuda_stream


def compute_mu_lam_from_E_nu(ctx, n_particles, logE, y, mu, lam):
    """utils.py:350-362 on CUDA tensors (in place into mu, lam)."""
    for t in (logE, y, mu, lam):
        if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
            raise _native.MpmkError("compute_mu_lam_from_E_nu needs contiguous float32 CUDA tensors")
    ctx.call("mpmk_compute_mu_lam", int(n_particles), logE.data_ptr(), y.data_ptr(), mu.data_ptr(), lam.data_ptr(),
             _stream())


def compute_mass_from_vol_density(n_particles, density, vol, mass):
    """utils.py:366-373 (one fp32 multiply per particle; torch elementwise on the device)."""
    torch.mul(density[:n_particles], vol[:n_particles], out=mass[:n_particles])
