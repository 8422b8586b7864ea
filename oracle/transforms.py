"""TEST INFRASTRUCTURE ONLY (see oracle/mpm_oracle.py): numpy float32 restatement of the reference's coordinate
changes around the solver, function by function -- utils/transform_utils.py and the chain main.py applies per frame.
Parity unpinned (the reference ships no vectors for them); pinned here by round trips and float64 re-derivations."""
import numpy as np

f32 = np.float32


def world2grid(means3D, grid_extent):
    """utils/transform_utils.py:8-15"""
    m = means3D.astype(f32)
    pos_min, pos_max = m.min(axis=0), m.max(axis=0)
    pos_center = ((pos_min + pos_max) / f32(2.0)).astype(f32)
    scaling_modifier = f32(f32(grid_extent) / f32(2.0) / (pos_max - pos_min).max())
    out = (m - pos_center) * scaling_modifier + np.ones(3, f32) * f32(grid_extent) / f32(2.0)
    return out.astype(f32), pos_center, scaling_modifier


def grid2world(means3D, covs, scaling_modifier, pos_center, grid_extent):
    """utils/transform_utils.py:18-21"""
    s = f32(scaling_modifier)
    m = (means3D.astype(f32) - np.ones(3, f32) * f32(grid_extent) / f32(2.0)) / s + pos_center.astype(f32)
    return m.astype(f32), (covs.astype(f32) / (s * s)).reshape(-1, 6).astype(f32)


def rotation_matrix(degree, axis):
    """utils/transform_utils.py:23-41"""
    t = f32(degree) / f32(180.0) * f32(3.1415926)
    c, s = np.cos(t, dtype=f32), np.sin(t, dtype=f32)
    return {0: np.array([[1, 0, 0], [0, c, -s], [0, s, c]], f32), 1: np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]], f32),
            2: np.array([[c, -s, 0], [s, c, 0], [0, 0, 1]], f32)}[axis]


def mat_from_upper(u):
    """:64-76"""
    return u.reshape(-1, 6)[:, [0, 1, 2, 1, 3, 4, 2, 4, 5]].reshape(-1, 3, 3)


def upper_from_mat(m):
    """:78-87"""
    return m.reshape(-1, 9)[:, [0, 1, 2, 4, 5, 8]]


def apply_rotations(p, Rs):
    """:47-49, :89-92"""
    for R in Rs:
        p = (p @ R.T).astype(f32)
    return p


def apply_cov_rotations(u, Rs):
    """:52-55, :95-99"""
    C = mat_from_upper(u)
    for R in Rs:
        C = (R @ (C @ R.T).astype(f32)).astype(f32)
    return upper_from_mat(C)


def apply_inverse_rotations(p, Rs):
    """:106-115"""
    for R in Rs[::-1]:
        p = (p @ R).astype(f32)
    return p


def apply_inverse_cov_rotations(u, Rs):
    """:118-123"""
    C = mat_from_upper(u)
    for R in Rs[::-1]:
        Q = R.T
        C = (Q @ (C @ Q.T).astype(f32)).astype(f32)
    return upper_from_mat(C)


def render_inputs(means_world, covs_world, Rs, pos_center, scaling_modifier=1.0):
    """main.py:134-146: what render_frame hands to the rasteriser as means3D / cov3D_precomp"""
    s = f32(scaling_modifier)
    m = pos_center.astype(f32) + (means_world.astype(f32) - np.ones(3, f32)) / s      # :139-142 (undoshift2center111, undotransform2origin)
    m = apply_inverse_rotations(m.astype(f32), Rs)
    c = (covs_world.astype(f32) / (s * s)).astype(f32)                                  # :145
    return m, apply_inverse_cov_rotations(c, Rs)                                         # :146
