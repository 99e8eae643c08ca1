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
    """model/common.py:623-631 (clamped at 1e-10, float32 result)."""
    return (-10.0 * np.log10(np.maximum(mse, 1e-10))).astype(np.float32)


# ---------------------------------------------------------------------------------------------------------------------
# Host-side helpers the unchanged callers import from model.common (train.py:17, vis/render.py:12, evaluation/eval.py:13).
# Small numpy / scipy routines outside the hot path, written against the behaviour of the reference's functions of the
# same name (file:line in each docstring); tests/test_cpu_host.py compares them with the reference's on the same inputs.
# ---------------------------------------------------------------------------------------------------------------------
def backup(out_dir, config):
    """model/common.py:492-506: snapshot of the run's config and sources under out_dir/backup.  Files that do not exist in
    the working directory are skipped rather than raising (the drop-in may be run from outside the reference tree)."""
    import os
    import shutil
    dst = os.path.join(out_dir, 'backup')
    os.makedirs(dst, exist_ok=True)
    shutil.copyfile(config, os.path.join(dst, 'config.yaml'))
    for f in ('train.py', './configs/default.yaml'):
        if os.path.isfile(f):
            shutil.copy(f, dst)
    for base in ('./model', './dataloading'):
        if not os.path.isdir(base):
            continue
        sub = os.path.join(dst, base[2:])
        os.makedirs(sub, exist_ok=True)
        for f in os.listdir(base):
            if os.path.isfile(os.path.join(base, f)):
                shutil.copy(os.path.join(base, f), sub)


def _slerp(times, rots_matrix):
    from scipy.spatial.transform import Rotation, Slerp
    return Slerp(times, Rotation.from_matrix(np.asarray(rots_matrix)))


def _poses_from(rots, trans):
    """rots (N,3,3) float32 tensor, trans (N,3,1) tensor -> (N,4,4)."""
    return convert3x4_4x4(torch.cat([rots, trans], dim=2))


def interp_poses(c2ws, N_views):
    """model/common.py:511-522: rotations by spherical interpolation over equally spaced keys, translations by linear
    interpolation (F.interpolate mode 'linear', align_corners False).  c2ws (N,4,4) cpu tensor -> (N_views,4,4)."""
    n = c2ws.shape[0]
    sl = _slerp(np.linspace(0, 1, n), c2ws[:, :3, :3])
    rots = torch.tensor(sl(np.linspace(0, 1, N_views)).as_matrix().astype(np.float32))
    trans = torch.nn.functional.interpolate(c2ws[:, :3, 3:].permute(2, 1, 0), size=N_views, mode='linear').permute(2, 1, 0)
    return _poses_from(rots, trans)


def scipy_bspline(cv, n=100, degree=3, periodic=False):
    """model/common.py:563-589: n samples of the clamped (or periodic) B-spline with control vertices cv."""
    import scipy.interpolate as si
    cv = np.asarray(cv)
    count = cv.shape[0]
    if periodic:
        kv = np.arange(-degree, count + degree + 1)
        factor, fraction = divmod(count + degree + 1, count)
        cv = np.roll(np.concatenate((cv,) * factor + (cv[:fraction],)), -1, axis=0)
        degree = np.clip(degree, 1, degree)
    else:
        degree = np.clip(degree, 1, count - 1)
        kv = np.clip(np.arange(count + degree + 1) - degree, 0, count - degree)
    max_param = count - (degree * (1 - periodic))
    return si.BSpline(kv, cv, degree)(np.linspace(0, max_param, n))


def interp_poses_bspline(c2ws, N_novel_imgs, input_times, degree):
    """model/common.py:523-531: translations on a B-spline through the camera centres, rotations by slerp over input_times."""
    trans = torch.tensor(scipy_bspline(c2ws[:, :3, 3], n=N_novel_imgs, degree=degree, periodic=False).astype(np.float32)).unsqueeze(2)
    sl = _slerp(input_times, c2ws[:, :3, :3])
    times = np.linspace(input_times[0], input_times[-1], N_novel_imgs)
    return _poses_from(torch.tensor(sl(times).as_matrix().astype(np.float32)), trans)


def interp_t(trans, input_times, target_times):
    """model/common.py:544-558 (its weights as written there: each neighbour is weighted by the distance to the OTHER side
    swapped, and a target that coincides with a key divides 0 by 0 -- kept, callers rely on neither)."""
    out = []
    for tt in target_times:
        diff = tt - input_times
        a = diff.copy()
        a[diff < 0] = 1000
        b = diff.copy()
        b[diff > 0] = -1000
        i1, i2 = np.argmin(a), np.argmin(-b)
        den = input_times[i2] - input_times[i1]
        out.append((tt - input_times[i1]) / den * trans[i1] + (input_times[i2] - tt) / den * trans[i2])
    return torch.stack(out, axis=0)


def get_poses_at_times(c2ws, input_times, target_times):
    """model/common.py:533-543."""
    sl = _slerp(input_times, c2ws[:, :3, :3])
    rots = torch.tensor(sl(target_times).as_matrix().astype(np.float32))
    return _poses_from(rots, interp_t(c2ws[:, :3, 3:], input_times, target_times))


def normalize(v):
    """model/common.py:371-373."""
    return v / np.linalg.norm(v)


def viewmatrix(z, up, pos):
    """model/common.py:374-380: camera frame (3,4) with forward axis z, approximate up vector and position."""
    f = normalize(z)
    right = normalize(np.cross(up, f))
    return np.stack([right, normalize(np.cross(f, right)), f, pos], 1)


def poses_avg(poses):
    """model/common.py:393-403: average pose of (N,3,5) [R | t | hwf] arrays."""
    hwf = poses[0, :3, -1:]
    frame = viewmatrix(normalize(poses[:, :3, 2].sum(0)), poses[:, :3, 1].sum(0), poses[:, :3, 3].mean(0))
    return np.concatenate([frame, hwf], 1)


def render_path_spiral(c2w, up, rads, focal, zdelta, zrate, rots, N):
    """model/common.py:381-392: N poses on a spiral around c2w looking at a point `focal` in front of it."""
    rads = np.array(list(rads) + [1.])
    hwf = c2w[:, 4:5]
    look_at = np.dot(c2w[:3, :4], np.array([0, 0, -focal, 1.]))
    out = []
    for theta in np.linspace(0., 2. * np.pi * rots, N + 1)[:-1]:
        c = np.dot(c2w[:3, :4], np.array([0.2 * np.cos(theta), -0.2 * np.sin(theta), -np.sin(theta * zrate) * 0.1, 1.]) * rads)
        out.append(np.concatenate([viewmatrix(normalize(c - look_at), up, c), hwf], 1))
    return out


def generate_spiral_nerf(learned_poses, bds, N_novel_views, hwf):
    """model/common.py:591-615: spiral of novel views around the average learned pose -> (N,3,4) tensor."""
    poses = np.concatenate((learned_poses[:, :3, :4].detach().cpu().numpy(), hwf[:len(learned_poses)]), axis=-1)
    c2w = poses_avg(poses)
    print('recentered', c2w.shape)
    up = normalize(poses[:, :3, 1].sum(0))
    close_depth, inf_depth = bds.min() * .9, bds.max() * 5.
    dt = .75
    focal = 1. / (((1. - dt) / close_depth + dt / inf_depth))
    rads = np.percentile(np.abs(poses[:, :3, 3]), 90, 0)
    path = render_path_spiral(c2w, up, rads, focal, close_depth * .2, zrate=.5, rots=2, N=N_novel_views)
    return torch.tensor(np.stack(path).astype(np.float32))[:, :3, :4]


def create_spheric_poses(radius, mean_h, n_poses=120):
    """model/common.py:333-369: circular path around the z axis -> (n_poses,3,4)."""
    def pose(theta, phi, r):
        trans = np.array([[1, 0, 0, 0], [0, 1, 0, 2 * mean_h], [0, 0, 1, -r]])
        rp = np.array([[1, 0, 0], [0, np.cos(phi), -np.sin(phi)], [0, np.sin(phi), np.cos(phi)]])
        rt = np.array([[np.cos(theta), 0, -np.sin(theta)], [0, 1, 0], [np.sin(theta), 0, np.cos(theta)]])
        return np.array([[-1, 0, 0], [0, 0, 1], [0, 1, 0]]) @ (rt @ rp @ trans)
    return np.stack([pose(th, -np.pi / 12, radius) for th in np.linspace(0, 2 * np.pi, n_poses + 1)[:-1]], 0)


def convert2mip(pts):
    """model/common.py:616-622: points outside the unit ball contracted into the ball of radius 2."""
    nrm = torch.linalg.norm(pts, ord=2, dim=-1)
    outside = nrm >= 1.0
    out = pts.clone()
    out[outside, :] = (2 - 1.0 / nrm[outside, None]) * (pts[outside, :] / nrm[outside, None])
    return out


def compute_errors(gt, pred):
    """model/common.py:676-697: depth metrics (abs_rel, sq_rel, rmse, rmse_log, a1, a2, a3) of two numpy arrays."""
    ratio = np.maximum(gt / pred, pred / gt)
    a1, a2, a3 = [(ratio < 1.25 ** k).mean() for k in (1, 2, 3)]
    rmse = np.sqrt(((gt - pred) ** 2).mean())
    rmse_log = np.sqrt(((np.log(gt) - np.log(pred)) ** 2).mean())
    return np.mean(np.abs(gt - pred) / gt), np.mean(((gt - pred) ** 2) / gt), rmse, rmse_log, a1, a2, a3


def to_pytorch(tensor, return_type=False):
    """model/common.py:42-57: numpy or torch -> a torch COPY (callers mutate the result in place)."""
    is_numpy = isinstance(tensor, np.ndarray)
    t = (torch.from_numpy(tensor) if is_numpy else tensor).clone()
    return (t, is_numpy) if return_type else t


def get_tensor_values(tensor, p, mode='nearest', scale=True, detach=True, detach_p=True, align_corners=False):
    """model/common.py:75-109: values of a (B,C,H,W) tensor at positions p (B,N,2) through grid_sample.  `scale` rescales p
    IN PLACE like the reference.  Used by the extractors on the host side of the render loop, not by the training step
    (which samples through nnb_gather_rays / nnb_reproj)."""
    _, _, h, w = tensor.shape
    if detach_p:
        p = p.detach()
    if scale:
        p[:, :, 0] = 2. * p[:, :, 0] / w - 1
        p[:, :, 1] = 2. * p[:, :, 1] / h - 1
    v = torch.nn.functional.grid_sample(tensor, p.unsqueeze(1), mode=mode, align_corners=align_corners).squeeze(2)
    if detach:
        v = v.detach()
    return v.permute(0, 2, 1)


def transform_to_camera_space(p_world, camera_mat, world_mat, scale_mat):
    """model/common.py:163-183: world points (B,N,3) -> camera space."""
    b, n, _ = p_world.shape
    ph = torch.cat([p_world, torch.ones(b, n, 1, device=p_world.device)], dim=-1).permute(0, 2, 1)
    return (camera_mat @ world_mat @ scale_mat @ ph)[:, :3].permute(0, 2, 1)


def project_to_cam(points, camera_mat, device):
    """model/common.py:436-457: (B,N,3) camera-space points -> normalised image coordinates and their in-frame mask."""
    B, N, _ = points.size()
    ph = torch.cat([points.permute(0, 2, 1), torch.ones(B, 1, N, device=device)], dim=1)
    xy = (camera_mat @ ph)[:, :3].permute(0, 2, 1)
    xy = xy[..., :2] / xy[..., 2:]
    return xy, (xy.abs().max(dim=-1)[0] <= 1).unsqueeze(-1).bool()


def normalize_tensor(tensor, min_norm=1e-5, feat_dim=-1):
    """model/common.py:265-276."""
    return tensor / torch.clamp(torch.norm(tensor, dim=feat_dim, keepdim=True), min=min_norm)


def vec2skew(v):
    """model/common.py:277-288: (3,) -> (3,3) skew-symmetric matrix."""
    z = torch.zeros((), dtype=v.dtype, device=v.device)
    return torch.stack([torch.stack([z, -v[2], v[1]]), torch.stack([v[2], z, -v[0]]), torch.stack([-v[1], v[0], z])])


def get_ndc_rays_fxfy(fxfy, near, rays_o, rays_d):
    """model/common.py:632-675, kept under its name for callers outside the renderer: rays (N,3),(N,3) with PER-RAY origins
    -> NDC origins, directions, as tensor expressions on the inputs' device.  The Renderer does not come through here: its
    rays share one origin and go through nnb_ndc_rays_fwd/bwd (_ops.NdcRays)."""
    t = -(near + rays_o[..., 2]) / rays_d[..., 2]
    o = rays_o + t[..., None] * rays_d
    ox, oy = o[..., 0] / o[..., 2], o[..., 1] / o[..., 2]
    o2 = 1. + 2. * near / o[..., 2]
    d0 = -1. / (1 / fxfy[0]) * (rays_d[..., 0] / rays_d[..., 2] - ox)
    d1 = -1. / (1 / fxfy[1]) * (rays_d[..., 1] / rays_d[..., 2] - oy)
    return (torch.stack([-1. / (1 / fxfy[0]) * ox, -1. / (1 / fxfy[1]) * oy, o2], -1), torch.stack([d0, d1, 1 - o2], -1))


# -- helpers the reference's own modules call internally; kept for callers that reach for them by name ---------------------
def skew_symmetric(w):
    """(...,3) -> (...,3,3) cross-product matrices (model/common.py:459-465)."""
    x, y, z = w.unbind(dim=-1)
    o = torch.zeros_like(x)
    return torch.stack([o, -z, y, z, o, -x, -y, x, o], dim=-1).reshape(w.shape[:-1] + (3, 3))


def _alternating_series(x, first, nth):
    """sum_i (-1)^i x^(2i) / d_i with d_0 = first and d_i = d_{i-1} (2i + k)(2i + k + 1), k fixed by `first`
    (1 -> sin x / x, 2 -> (1 - cos x) / x^2, 6 -> (x - sin x) / x^3): model/common.py:467-490."""
    k = {1.0: 0, 2.0: 1, 6.0: 2}[float(first)]
    total, denom, x2 = torch.zeros_like(x), float(first), x * x
    power = torch.ones_like(x)
    for i in range(nth + 1):
        if i > 0:
            denom *= (2 * i + k) * (2 * i + k + 1)
            power = power * x2
        total = total + (-1) ** i * power / denom
    return total


def taylor_A(x, nth=10):
    return _alternating_series(x, 1.0, nth)


def taylor_B(x, nth=10):
    return _alternating_series(x, 2.0, nth)


def taylor_C(x, nth=10):
    return _alternating_series(x, 6.0, nth)


def check_weights(params):
    """Warn about NaNs in a state dict (model/common.py:240-248)."""
    import logging
    for name, value in params.items():
        if torch.isnan(value).any():
            logging.getLogger(__name__).warning('NaN Values detected in model weight %s.' % name)


def check_tensor(tensor, tensorname='', input_tensor=None):
    """Warn about NaNs in a tensor (model/common.py:251-262)."""
    import logging
    if torch.isnan(tensor).any():
        log = logging.getLogger(__name__)
        log.warning('Tensor %s contains nan values.' % tensorname)
        if input_tensor is not None:
            log.warning('Input was: %s' % (input_tensor,))


def reprojection(pixels, depth, Rt_ref, world_mat, camera_mat):
    """Pixels (B,N,2) with depths of one view -> their image coordinates in a reference view, plus the inside-the-image mask
    (model/common.py:405-435): K Rt_ref W^-1 K^-1 [x d, y d, d, 1]^T, then the perspective divide."""
    assert pixels.shape[-1] == 2
    pixels, was_numpy = to_pytorch(pixels, True)
    depth, camera_mat, Rt_ref = to_pytorch(depth), to_pytorch(camera_mat), to_pytorch(Rt_ref)
    d = depth.reshape(1, -1, 1)
    ones = torch.ones_like(pixels)
    homog = torch.cat([pixels * d, d.expand(pixels.shape[0], -1, 1), ones[..., :1]], dim=-1).transpose(1, 2)   # (B,4,N)
    chain = camera_mat @ Rt_ref @ torch.inverse(world_mat) @ torch.inverse(camera_mat)
    cam = (chain @ homog)[:, :3].transpose(1, 2)
    xy = cam[..., :2] / cam[..., 2:]
    valid = (xy.abs().amax(dim=-1) <= 1).unsqueeze(-1).float()
    return (xy.numpy() if was_numpy else xy), valid
