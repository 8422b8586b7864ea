"""The per-frame render step of the reference's main.py (render_frame, main.py:108-157; camera set-up, main.py:84-106,
244-267) on the UNCHANGED diff-gaussian-rasterization the reference ships as a submodule.  Nothing of the rasteriser lives
in this repository: `baseline/build_rasterizer.py` compiles it for sm_100a from the reference tree into the git-ignored
`baseline/_ref/`, and this module only imports it (or any `diff_gaussian_rasterization` already importable).

The solver hands it `export_world(..., mode="render")`'s tensors as they are: `means3D [N,3]`, `cov3D_precomp [N,6]`, the
export kernel's persistent buffers, already through render_frame's undo chain (main.py:139-146).  The rasteriser launches
on the legacy default stream (cuda_rasterizer/forward.cu:389,428): a solver running on another stream has to be waited
for first (`torch.cuda.current_stream().wait_stream(solver_stream)`)."""
import math
import os
import sys

import numpy as np
import torch

_REF = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref")


def load_rasterizer():
    """the module `diff_gaussian_rasterization` (GaussianRasterizationSettings, GaussianRasterizer); raises with the
    build recipe if it is neither importable nor built under baseline/_ref/"""
    try:
        import diff_gaussian_rasterization as dgr
        return dgr
    except ImportError:
        pass
    if os.path.isdir(os.path.join(_REF, "diff_gaussian_rasterization")):
        if _REF not in sys.path:
            sys.path.insert(0, _REF)
        import diff_gaussian_rasterization as dgr
        return dgr
    raise RuntimeError("diff_gaussian_rasterization is not built: run `python baseline/build_rasterizer.py` where the reference "
                       "tree is mounted (it compiles the reference's unchanged submodule for sm_100a into baseline/_ref/)")


class Camera:
    """what render_frame reads off its TinyCam (utils/render_utils.py): size, fields of view, the transposed world-to-view
    matrix, the transposed full projection and the camera centre"""

    def __init__(self, width, height, fov_x, fov_y, view_mat, full_proj_mat, cam_center):
        self.width, self.height, self.FovX, self.FovY = int(width), int(height), float(fov_x), float(fov_y)
        self.view_mat, self.full_proj_mat, self.cam_center = view_mat, full_proj_mat, cam_center


def _projection(znear, zfar, fov_x, fov_y):
    """the perspective matrix of gaussian_splatting/utils/graphics_utils.py:51-71 (z forward, depth in [0, 1])"""
    tx, ty = math.tan(fov_x / 2), math.tan(fov_y / 2)
    P = torch.zeros(4, 4)
    P[0, 0], P[1, 1] = 1.0 / tx, 1.0 / ty
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def orbit_camera(center, frame, device, azimuth=130.0, elevation=10.0, radius=5.75, width=800, height=800, fov_x=0.69, fov_y=0.69):
    """The camera main.py:84-106 puts on a sphere around `center` (modify_cam: azimuth 130 deg, elevation 10 deg, radius
    5.75): `frame` = [h1 | h2 | vertical] as columns (the observant coordinates of main.py:244-262); the camera looks at
    the centre with the frame's vertical axis up."""
    center = np.asarray(center, np.float64)
    frame = np.asarray(frame, np.float64)
    az, el = math.radians(azimuth), math.radians(elevation)
    eye = center + frame @ (radius * np.array([math.cos(az) * math.cos(el), math.sin(az) * math.cos(el), math.sin(el)]))
    fwd = center - eye
    fwd /= np.linalg.norm(fwd)
    down = -frame[:, 2] - np.dot(-frame[:, 2], fwd) * fwd          # image y points down
    down /= np.linalg.norm(down)
    right = np.cross(down, fwd)
    c2w = np.eye(4)
    c2w[:3, :3] = np.column_stack((right, down, fwd))
    c2w[:3, 3] = eye
    w2c = np.linalg.inv(c2w)
    view = torch.tensor(w2c, dtype=torch.float32).t().contiguous().to(device)
    proj = _projection(0.01, 100.0, fov_x, fov_y).t().contiguous().to(device)
    full = view.unsqueeze(0).bmm(proj.unsqueeze(0)).squeeze(0).contiguous()
    return Camera(width, height, fov_x, fov_y, view, full, torch.tensor(eye, dtype=torch.float32, device=device))


def render_frame(cam, means3D, cov3D, shs, opacities, bg_color, sh_degree=0, scaling_modifier=1.0, rasterizer=None):
    """main.py:108-157 after the undo chain (which export_world(mode="render") has already applied): rasterise the simulated
    Gaussians with precomputed 3-D covariances.  Returns the image as a CUDA tensor [3, H, W] (the caller decides when to
    bring it to the host) and the screen radii."""
    dgr = rasterizer or load_rasterizer()
    settings = dgr.GaussianRasterizationSettings(
        image_height=cam.height, image_width=cam.width, tanfovx=math.tan(cam.FovX * 0.5), tanfovy=math.tan(cam.FovY * 0.5),
        bg=bg_color, scale_modifier=scaling_modifier, viewmatrix=cam.view_mat, projmatrix=cam.full_proj_mat,
        sh_degree=sh_degree, campos=cam.cam_center, prefiltered=False, debug=False)
    image, radii = dgr.GaussianRasterizer(raster_settings=settings)(
        means3D=means3D, means2D=None, shs=shs, colors_precomp=None, opacities=opacities, scales=None, rotations=None,
        cov3D_precomp=cov3D)
    return image, radii


def to8b(image):
    """[3, H, W] float image -> [H, W, 3] uint8 (utils/render_utils.py to8b)"""
    return (255.0 * image.detach().clamp(0.0, 1.0)).to(torch.uint8).permute(1, 2, 0).contiguous().cpu().numpy()


def save_png(image, path):
    from PIL import Image
    os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
    Image.fromarray(to8b(image)).save(path)
