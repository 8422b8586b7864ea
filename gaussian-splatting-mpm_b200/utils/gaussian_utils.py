"""What main.py needs from GaussianModel on the way into the solver: activations and the 3D covariance of every
Gaussian (gaussian_splatting/scene/gaussian_model.py:27-41,117-118; utils/general_utils.py:64-110), device-agnostic."""
import torch


def build_rotation(r):
    """general_utils.py:78-99: unit quaternion (w, x, y, z) -> rotation matrix"""
    q = r / torch.sqrt((r * r).sum(dim=1))[:, None]
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                     2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                     2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], dim=1)
    return R.reshape(-1, 3, 3)


def strip_symmetric(sym):
    """general_utils.py:64-76: upper triangle xx, xy, xz, yy, yz, zz"""
    return sym.reshape(-1, 9)[:, [0, 1, 2, 4, 5, 8]]


def covariance_from_scaling_rotation(log_scaling, rotation, scaling_modifier=1.0):
    """get_covariance(): scaling = exp(_scaling) (:33), L = R diag(s * modifier), Sigma = L L^T (:27-31)"""
    s = torch.exp(log_scaling) * scaling_modifier
    L = build_rotation(rotation) * s[:, None, :]
    return strip_symmetric(L @ L.transpose(1, 2))


def opacity_activation(raw):
    return torch.sigmoid(raw)
