"""Geometry helpers with the reference's names (model/common.py), backed by the stage-1 kernels."""
import numpy as np
import torch

from . import _ops


def arange_pixels(resolution=(128, 128), batch_size=1, image_range=(-1., 1.), device=torch.device("cuda")):
    """model/common.py:13-40 -> (int64 locations (B,h*w,2), fp32 scaled (B,h*w,2))."""
    if tuple(image_range) != (-1., 1.):
        raise NotImplementedError("image_range other than (-1, 1) is never used by the reference")
    h, w = resolution
    loc, scaled = _ops.pixel_grid(h, w, device)
    return loc.unsqueeze(0).repeat(batch_size, 1, 1), scaled.unsqueeze(0).repeat(batch_size, 1, 1)


def get_mask(tensor):
    """model/common.py:60-72: mask of finite values."""
    if isinstance(tensor, np.ndarray):
        return np.isfinite(tensor)
    return torch.isfinite(tensor)


def make_c2w(r, t):
    """model/common.py:301-310 via the SE(3) exp-map kernel.  r, t: (3,)"""
    return _ops.PoseC2W.apply(r.reshape(1, 3), t.reshape(1, 3), None, 0)


def Exp(r):
    """model/common.py:290-299."""
    return make_c2w(r, torch.zeros_like(r))[:3, :3]


def convert3x4_4x4(input):
    """model/common.py:312-330 (torch tensors only)."""
    if input.dim() == 3:
        out = torch.cat([input, torch.zeros_like(input[:, 0:1])], dim=1)
        out[:, 3, 3] = 1.0
        return out
    return torch.cat([input, torch.tensor([[0, 0, 0, 1]], dtype=input.dtype, device=input.device)], dim=0)


def origin_to_world(n_points, camera_mat, world_mat, scale_mat, invert=True):
    """model/common.py:186-215 -> (1,n_points,3)."""
    if not invert:
        raise NotImplementedError("invert=False is never used on the hot path")
    pix = torch.zeros(1, 1, 2, device=camera_mat.device)
    origin = _ops.RayGen.apply(pix, None, camera_mat, world_mat, scale_mat, True)[0]
    return origin.reshape(1, 1, 3).expand(1, n_points, 3)


def transform_to_world(pixels, depth, camera_mat, world_mat=None, scale_mat=None, invert=True, device=None):
    """model/common.py:112-160 -> (1,N,3) = origin + unnormalised ray * depth."""
    if not invert:
        raise NotImplementedError("invert=False is never used on the hot path")
    eye = torch.eye(4, device=pixels.device).unsqueeze(0)
    world_mat = eye if world_mat is None else world_mat
    scale_mat = eye if scale_mat is None else scale_mat
    origin, dirs, nrm, _, _ = _ops.RayGen.apply(pixels, depth, camera_mat, world_mat, scale_mat, False)
    return (origin.reshape(1, 1, 3) + dirs.unsqueeze(0) * depth.reshape(1, -1, 1))


def image_points_to_world(image_points, camera_mat, world_mat, scale_mat, invert=True):
    """model/common.py:218-237."""
    d = torch.ones(1, image_points.shape[1], 1, device=image_points.device)
    return transform_to_world(image_points, d, camera_mat, world_mat, scale_mat, invert=invert)


def mse2psnr(mse):
    """model/common.py:623."""
    return -10.0 * np.log10(mse)
