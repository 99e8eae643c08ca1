"""Geometry path of the reference's Renderer: sphere tracing of the occupancy field and phong shading.

  get_sphere_intersection   model/rendering.py:440-460
  ray_marching              model/rendering.py:274-385   (512 proposals per ray, first outside->inside sign change)
  secant                    model/rendering.py:387-419   (8 refinement steps)
  phong_renderer            model/rendering.py:203-271   (normals = -d raw_density / d p, light at the camera)

The field evaluations (occupancy at explicit points, and its input gradient for the normals) are C-ABI kernels
(`nnb_mlp_points_fwd`, `nnb_mlp_points_density_grad`); the bookkeeping between them (proposal depths, sign-change search,
interval updates, shading) is elementwise tensor math on the device with no host synchronisation except the two data-dependent
compactions the reference also performs (rays with a sign change, surface points).  Everything is forward only, as in the
reference (`torch.no_grad()` / `.detach()` throughout)."""
import numpy as np
import torch

from . import _ops


def get_sphere_intersection(cam_loc, ray_directions, r=1.0):
    """cam_loc (3,), ray_directions (R,3) unit -> (R,2) near/far distances clamped at 0 (zeros where the ray misses), mask (R,)."""
    ray_cam_dot = ray_directions @ cam_loc
    under_sqrt = ray_cam_dot ** 2 - (cam_loc.norm(2) ** 2 - r ** 2)
    mask = under_sqrt > 0
    root = torch.sqrt(under_sqrt.clamp_min(0))
    inter = torch.stack([-root - ray_cam_dot, root - ray_cam_dot], -1)
    inter = torch.where(mask.unsqueeze(-1), inter, torch.zeros_like(inter)).clamp_min(0.0)
    return inter, mask


def _occupancy(model, pts, tau, max_points):
    out = []
    for p in torch.split(pts, int(max_points), dim=0):
        out.append(model(p, only_occupancy=True).reshape(-1) - tau)
    return torch.cat(out, 0)


def secant(model, f_low, f_high, d_low, d_high, n_secant_steps, ray0, ray_dirs, tau):
    """model/rendering.py:387-419 on the rays that bracket a surface.  ray0 (3,) shared or (M,3); ray_dirs (M,3)."""
    d_pred = -f_low * (d_high - d_low) / (f_high - f_low) + d_low
    for _ in range(n_secant_steps):
        p_mid = ray0 + d_pred.unsqueeze(-1) * ray_dirs
        f_mid = model(p_mid, only_occupancy=True).reshape(-1) - tau
        low = f_mid < 0
        d_low = torch.where(low, d_pred, d_low)
        f_low = torch.where(low, f_mid, f_low)
        d_high = torch.where(low, d_high, d_pred)
        f_high = torch.where(low, f_high, f_mid)
        d_pred = -f_low * (d_high - d_low) / (f_high - f_low) + d_low
    return d_pred


def _proposals_fused(renderer, cam, ray_dirs, d_prop, tau):
    """Occupancy at the (R, n) proposal depths through the fused tensor-core render kernel: every run of 128 consecutive
    proposals of a ray is one pseudo-ray with its own origin cam + dir * d[128 j] and direction dir * (d[128 j + 127] - d[128 j]),
    sampled at linspace(0, 1, 128) without jitter; the kernel's per-sample `alpha` output is the occupancy (dist_alpha off)."""
    R, n = d_prop.shape
    seg = n // 128
    d0 = d_prop[:, 0::128][:, :seg]                                   # (R, seg) first depth of each run
    d1 = d_prop[:, 127::128][:, :seg]
    ray_o = (cam.view(1, 1, 3) + ray_dirs.unsqueeze(1) * d0.unsqueeze(-1)).reshape(-1, 3).contiguous()
    dirs = (ray_dirs.unsqueeze(1) * (d1 - d0).unsqueeze(-1)).reshape(-1, 3).contiguous()
    model = renderer.model
    params = [p.detach() for p in model.param_list()]
    packed = renderer._packed_weights(model.param_list())
    rc = _ops.make_cfg(R * seg, 128, 0.0, 1.0, model.occ_activation == 'softplus', 0, 0, 0, renderer._prec())
    _, _, _, alpha, _ = _ops.render_nograd(rc, params, packed, cam, dirs, None, want_aux='alpha', ray_o=ray_o, view=dirs)
    return alpha.view(R, n) - tau


def ray_marching(model, cam, ray_dirs, tau=0.5, n_steps=512, n_secant_steps=8, depth_range=(0., 2.4), max_points=3500000, rad=1.0,
                 renderer=None):
    """cam (3,), ray_dirs (R,3) -> d (R,): surface distance, inf where no outside->inside crossing, 0 where the first sample is
    already occupied (model/rendering.py:274-385).  With `renderer` on the bf16 path and n_steps a multiple of 128 the proposals
    are evaluated by the fused tensor-core kernel (the secant refinement stays on the fp32 point kernels)."""
    R = ray_dirs.shape[0]
    dev = ray_dirs.device
    inter, _ = get_sphere_intersection(cam, ray_dirs, r=rad)
    d_far = inter[:, 1]
    u = torch.linspace(0, 1, steps=n_steps, device=dev).view(1, n_steps)
    d_prop = depth_range[0] * (1. - u) + d_far.view(R, 1) * u                      # (R, n)
    if renderer is not None and renderer.precision == 'bf16' and n_steps % 128 == 0 and not model.dist_alpha:
        val = _proposals_fused(renderer, cam, ray_dirs, d_prop, tau)
    else:
        # rows of rays per occupancy call so that one call stays below max_points
        rows = max(1, int(max_points) // n_steps)
        vals = []
        for a in range(0, R, rows):
            p = cam.view(1, 1, 3) + ray_dirs[a:a + rows].unsqueeze(1) * d_prop[a:a + rows].unsqueeze(-1)
            vals.append(_occupancy(model, p.reshape(-1, 3), tau, max_points).view(-1, n_steps))
        val = torch.cat(vals, 0)
    mask_0_not_occupied = val[:, 0] < 0
    sign_matrix = torch.cat([torch.sign(val[:, :-1] * val[:, 1:]), torch.ones(R, 1, device=dev)], dim=-1)
    cost_matrix = sign_matrix * torch.arange(n_steps, 0, -1, device=dev).float()
    values, indices = torch.min(cost_matrix, -1)
    mask_sign_change = values < 0
    ar = torch.arange(R, device=dev)
    mask_neg_to_pos = val[ar, indices] < 0
    mask = mask_sign_change & mask_neg_to_pos & mask_0_not_occupied
    hi = torch.clamp(indices + 1, max=n_steps - 1)
    d_low, f_low = d_prop[ar, indices][mask], val[ar, indices][mask]
    d_high, f_high = d_prop[ar, hi][mask], val[ar, hi][mask]
    d_out = torch.ones(R, device=dev)
    if d_low.shape[0] != 0:
        d_out[mask] = secant(model, f_low, f_high, d_low, d_high, n_secant_steps, cam.view(1, 3), ray_dirs[mask], tau)
    d_out[~mask] = np.inf
    d_out[~mask_0_not_occupied] = 0
    return d_out


def phong_render(renderer, origin, ray_dirs, it=0):
    """model/rendering.py:203-271 from the camera centre `origin` (3,) and unit rays (R,3).
    -> dict rgb (1,R,3) shaded, normal None, rgb_surf (1,R,3) network colour at the surface."""
    model = renderer.model
    dev = ray_dirs.device
    rad = renderer.cfg['radius']
    light = (origin / origin.norm(2)).view(3, 1)
    diffuse_per = torch.tensor([0.7, 0.7, 0.7], device=dev)
    ambient = torch.tensor([0.3, 0.3, 0.3], device=dev)
    was_training = model.training
    model.eval()
    with torch.no_grad():
        d_i = ray_marching(model, origin, ray_dirs, n_secant_steps=8, n_steps=512, rad=rad, renderer=renderer)
        mask_zero_occupied = d_i == 0
        mask_pred = torch.isfinite(d_i)
        dists = torch.ones_like(d_i)
        dists[mask_pred] = d_i[mask_pred]
        dists[mask_zero_occupied] = 0.
        obj = mask_pred & ~mask_zero_occupied
        points = origin.view(1, 3) + ray_dirs * dists.unsqueeze(-1)
        view = -ray_dirs
        rgb_values = torch.ones_like(points)
        rgb_val = torch.zeros_like(points)
        surf, surf_view = points[obj], view[obj]
        if surf.shape[0] > 0:
            grad = torch.cat([model.gradient(p, it)[:, 0, :] for p in torch.split(surf, 1000000, dim=0)], 0)
            normals = grad / grad.norm(2, 1, keepdim=True)
            diffuse = torch.mm(normals, light).clamp_min(0).repeat(1, 3) * diffuse_per.unsqueeze(0)
            rgb_values[obj] = (ambient.unsqueeze(0) + diffuse).clamp_max(1.0)
            rgb_val[obj] = model(surf, surf_view)
    if was_training:
        model.train()
    return {'rgb': rgb_values.reshape(1, -1, 3), 'normal': None, 'rgb_surf': rgb_val.reshape(1, -1, 3)}
