"""Drop-in for the reference's utils/transform_utils.py (the coordinate changes either side of the MPM solver,
main.py:189-242 and :135-146, :310-313).  Same names, arguments and arithmetic; tensors stay on the device they
come in on (the reference hard-codes .cuda()).  The per-frame chain grid2world -> undo transforms is also available
fused into the solver's export kernel: MPM_Simulator.export_world() (csrc/export.cu k_export_world)."""
import numpy as np
import torch


def world2grid(means3D, sim_args):
    """utils/transform_utils.py:8-15: longest axis of the bounding box -> grid_extent/2, centred in the grid"""
    pos_min, pos_max = means3D.min(dim=0)[0], means3D.max(dim=0)[0]
    pos_center = (pos_min + pos_max) / 2.0
    scaling_modifier = sim_args.grid_extent / 2.0 / (pos_max - pos_min).max()
    transformed_means3D = (means3D - pos_center) * scaling_modifier + torch.ones(3, device=means3D.device) * sim_args.grid_extent / 2.0
    return transformed_means3D, pos_center, scaling_modifier


def grid2world(means3D, covs, scaling_modifier, pos_center, sim_args):
    """utils/transform_utils.py:18-21"""
    transformed_means3D = (means3D - torch.ones(3, device=means3D.device) * sim_args.grid_extent / 2.0) / scaling_modifier + pos_center
    transformed_covs = covs / (scaling_modifier * scaling_modifier)
    return transformed_means3D, transformed_covs.view(-1, 6)


def generate_rotation_matrix(degree, axis, device=None):
    """utils/transform_utils.py:23-41 (pi truncated to 3.1415926 as there)"""
    degree = torch.as_tensor(degree, dtype=torch.float32)
    cos_theta = torch.cos(degree / 180.0 * 3.1415926)
    sin_theta = torch.sin(degree / 180.0 * 3.1415926)
    if axis == 0:
        rotation_matrix = torch.tensor([[1, 0, 0], [0, cos_theta, -sin_theta], [0, sin_theta, cos_theta]])
    elif axis == 1:
        rotation_matrix = torch.tensor([[cos_theta, 0, sin_theta], [0, 1, 0], [-sin_theta, 0, cos_theta]])
    elif axis == 2:
        rotation_matrix = torch.tensor([[cos_theta, -sin_theta, 0], [sin_theta, cos_theta, 0], [0, 0, 1]])
    else:
        raise ValueError("Invalid axis selection")
    if device is None:
        device = "cuda" if torch.cuda.is_available() else "cpu"
    return rotation_matrix.to(device)


def generate_rotation_matrices(degrees, axises, device=None):
    assert len(degrees) == len(axises)
    return [generate_rotation_matrix(degrees[i], axises[i], device) for i in range(len(degrees))]


def apply_rotation(position_tensor, rotation_matrix):
    return torch.mm(position_tensor, rotation_matrix.T)


def apply_cov_rotation(cov_tensor, rotation_matrix):
    rotated = torch.matmul(cov_tensor, rotation_matrix.T)
    return torch.matmul(rotation_matrix, rotated)


def get_mat_from_upper(upper_mat):
    upper_mat = upper_mat.reshape(-1, 6)
    return upper_mat[:, [0, 1, 2, 1, 3, 4, 2, 4, 5]].reshape(-1, 3, 3)


def get_upper_from_mat(mat):
    return mat.reshape(-1, 9)[:, [0, 1, 2, 4, 5, 8]]


def apply_rotations(position_tensor, rotation_matrices):
    for R in rotation_matrices:
        position_tensor = apply_rotation(position_tensor, R)
    return position_tensor


def apply_cov_rotations(upper_cov_tensor, rotation_matrices):
    cov_tensor = get_mat_from_upper(upper_cov_tensor)
    for R in rotation_matrices:
        cov_tensor = apply_cov_rotation(cov_tensor, R)
    return get_upper_from_mat(cov_tensor)


def undoshift2center111(position_tensor):
    return position_tensor - torch.tensor([1.0, 1.0, 1.0], device=position_tensor.device)


def apply_inverse_rotation(position_tensor, rotation_matrix):
    return torch.mm(position_tensor, rotation_matrix)


def apply_inverse_rotations(position_tensor, rotation_matrices):
    for i in range(len(rotation_matrices)):
        position_tensor = apply_inverse_rotation(position_tensor, rotation_matrices[len(rotation_matrices) - 1 - i])
    return position_tensor


def apply_inverse_cov_rotations(upper_cov_tensor, rotation_matrices):
    cov_tensor = get_mat_from_upper(upper_cov_tensor)
    for i in range(len(rotation_matrices)):
        R = rotation_matrices[len(rotation_matrices) - 1 - i]
        cov_tensor = apply_cov_rotation(cov_tensor, R.T)
    return get_upper_from_mat(cov_tensor)


def undotransform2origin(position_tensor, scale, original_mean_pos):
    return original_mean_pos + position_tensor / scale


def undo_all_transforms(input, rotation_matrices, scale_origin, original_mean_pos):
    return apply_inverse_rotations(undotransform2origin(undoshift2center111(input), scale_origin, original_mean_pos), rotation_matrices)


def simulatable_mask(positions, sim_area):
    """main.py:189-194: Gaussians inside the axis-aligned box sim_area = [min xyz, max xyz]"""
    bound = torch.as_tensor(np.array(sim_area), dtype=positions.dtype, device=positions.device)
    return torch.logical_and((positions <= bound[1]).all(dim=1), (positions >= bound[0]).all(dim=1))


# ---- the composite maps export_world() hands to the kernel ----------------------------------------------------
def inverse_rotation_composite(rotation_matrices, device=None):
    """M with apply_inverse_rotations(p, Rs) == p @ M and apply_inverse_cov_rotations(C, Rs) == M^T C M (float64)"""
    M = np.eye(3)
    for i in range(len(rotation_matrices)):
        R = rotation_matrices[len(rotation_matrices) - 1 - i]
        M = M @ (R.detach().cpu().double().numpy() if torch.is_tensor(R) else np.asarray(R, np.float64))
    return M


def world_affine(mode, scaling_modifier, pos_center, grid_extent, render_scaling_modifier=1.0):
    """(alpha, beta[3], gamma): position = x * alpha + beta, covariance = cov * gamma, before the inverse rotations.
    mode "grid2world": main.py:310-313.  mode "render": that followed by render_frame's own undo chain
    (main.py:139-145: minus (1,1,1), / its scaling_modifier argument -- 1.0 at every call site -- plus pos_center)."""
    s = float(scaling_modifier)
    c = np.asarray(pos_center.detach().cpu().double().numpy() if torch.is_tensor(pos_center) else pos_center, np.float64)
    alpha, beta, gamma = 1.0 / s, c - 0.5 * float(grid_extent) / s, 1.0 / (s * s)
    if mode == "render":
        r = float(render_scaling_modifier)
        alpha, beta, gamma = alpha / r, (beta - 1.0) / r + c, gamma / (r * r)
    elif mode != "grid2world":
        raise ValueError(mode)
    return alpha, beta, gamma
