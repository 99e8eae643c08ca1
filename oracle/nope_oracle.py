"""CPU ORACLE -- TEST INFRASTRUCTURE ONLY.  Never imported by the product package.

A plain-PyTorch (CPU, fp32 or fp64) restatement of the NoPe-NeRF training / render hot
path of js0n-lai/my-nope-nerf, written stage by stage so every CUDA kernel has a
function to be compared with.  Only `tests/`, `__graft_entry__.smoke()` and the
`cpu_baseline` / `--impl reference` legs of `bench.py` may import this file.

Parity status: the reference ships NO golden vectors or tests, so
this oracle is pinned against outputs of the reference itself, generated in the build
container by `oracle/make_golden.py` (which imports /root/reference read-only) and
committed under `tests/golden/`.  `tests/test_oracle_golden.py` checks every function
here against those fixtures.

Each function cites the reference file:line it restates (paths relative to the
reference root).  Everything is functional: parameters come in as a flat dict with the
reference's `state_dict` key names (without the `renderer.model.` prefix).
"""
import math

import torch
import torch.nn.functional as F

EPS_T = 1e-6  # model/rendering.py:9

# ----------------------------------------------------------------------------------------------
# Stage 1: poses, pixel grid, ray generation
# ----------------------------------------------------------------------------------------------


def hat(v):
    """3-vector -> skew-symmetric matrix.  model/common.py:277-287 (vec2skew)."""
    z = v.new_zeros(())
    return torch.stack([
        torch.stack([z, -v[2], v[1]]),
        torch.stack([v[2], z, -v[0]]),
        torch.stack([-v[1], v[0], z]),
    ])


def exp_so3(r):
    """Rodrigues formula with the reference's `|r| + 1e-15` guard (no Taylor branch).
    model/common.py:290-299 (Exp)."""
    K = hat(r)
    th = r.norm() + 1e-15
    eye = torch.eye(3, dtype=r.dtype, device=r.device)
    return eye + (torch.sin(th) / th) * K + ((1 - torch.cos(th)) / th ** 2) * (K @ K)


def pose_c2w(r, t, init_c2w=None):
    """c2w = [Exp(r) | t; 0 0 0 1] @ init_c2w.  model/common.py:301-310, model/poses.py:23-31."""
    top = torch.cat([exp_so3(r), t.reshape(3, 1)], dim=1)
    bottom = torch.tensor([[0, 0, 0, 1]], dtype=r.dtype, device=r.device)
    c2w = torch.cat([top, bottom], dim=0)
    if init_c2w is not None:
        c2w = c2w @ init_c2w
    return c2w


def distortion(global_scales, global_shifts, cam_id, fix_scaleN=True):
    """model/distortions.py:19-27.  Below 0.01 the scale becomes a constant (gradient cut);
    the last camera's scale is pinned to 1 when fix_scaleN."""
    n = global_scales.shape[0]
    scale = global_scales[cam_id]
    if float(scale.detach()) < 0.01:
        scale = torch.tensor(0.01, dtype=global_scales.dtype)
    if fix_scaleN and int(cam_id) == n - 1:
        scale = torch.tensor(1, dtype=global_scales.dtype)
    return scale, global_shifts[cam_id]


def pixel_grid(h, w, dtype=torch.float32):
    """Scaled pixel centres in [-1,1], x fastest.  model/common.py:13-40 (arange_pixels).
    Returns (int64 locations (1,h*w,2), float scaled (1,h*w,2))."""
    ys, xs = torch.meshgrid(torch.arange(h), torch.arange(w), indexing="ij")
    loc = torch.stack([xs, ys], dim=-1).reshape(1, -1, 2).long()
    f = loc.to(dtype)
    px = 2.0 * f[..., 0] / (w - 1) - 1.0
    py = 2.0 * f[..., 1] / (h - 1) - 1.0
    return loc, torch.stack([px, py], dim=-1)


def unproject(pixels, depth, camera_mat, world_mat=None, scale_mat=None):
    """S^-1 W^-1 K^-1 [px*d, py*d, d, 1]^T.  model/common.py:112-160 (transform_to_world).
    pixels (1,R,2), depth (1,R,1), matrices (1,4,4) -> (1,R,3)."""
    eye = torch.eye(4, dtype=pixels.dtype).unsqueeze(0)
    world_mat = eye if world_mat is None else world_mat
    scale_mat = eye if scale_mat is None else scale_mat
    A = torch.inverse(scale_mat) @ torch.inverse(world_mat) @ torch.inverse(camera_mat)
    hom = torch.cat([pixels, torch.ones_like(pixels)], dim=-1)  # (1,R,4) = [px,py,1,1]
    hom = torch.cat([hom[..., :3] * depth, hom[..., 3:]], dim=-1)
    return (A @ hom.transpose(1, 2))[:, :3].transpose(1, 2)


def camera_origin(camera_mat, world_mat, scale_mat):
    """S^-1 W^-1 K^-1 [0,0,0,1]^T.  model/common.py:186-215 (origin_to_world). -> (1,3)"""
    A = torch.inverse(scale_mat) @ torch.inverse(world_mat) @ torch.inverse(camera_mat)
    return A[:, :3, 3]


def ray_setup(pixels, depth, camera_mat, world_mat, scale_mat, normalise_ray=True):
    """Ray origin / direction / supervision distance / validity mask.
    model/rendering.py:53-80.  Returns origin (3,), dirs (R,3), dir_norm (R,), d_gt (R,), mask (R,) bool."""
    o = camera_origin(camera_mat, world_mat, scale_mat)  # (1,3)
    surf = unproject(pixels, depth, camera_mat, world_mat, scale_mat)
    d_gt = torch.norm(surf - o.unsqueeze(1), p=2, dim=-1)  # (1,R)
    plane = unproject(pixels, torch.ones_like(depth), camera_mat, world_mat, scale_mat)
    d = plane - o.unsqueeze(1)
    nrm = d.norm(2, 2)
    if normalise_ray:
        d = d / nrm.unsqueeze(-1)
    else:
        d_gt = d_gt / nrm
    finite = torch.isfinite(d_gt)
    mask = finite & ~(d_gt == 0)
    return o[0], d[0], nrm[0], d_gt[0], mask[0]


# ----------------------------------------------------------------------------------------------
# Stage 2: sampling, positional encoding, MLP
# ----------------------------------------------------------------------------------------------


def stratified_z(n_rays, n_samples, near, far, noise=None, dtype=torch.float32):
    """model/rendering.py:89-90,183-191 (sample_uniform).  `noise` (n_rays, n_samples) in [0,1) or None."""
    u = torch.linspace(0.0, 1.0, steps=n_samples, dtype=dtype)
    nf = torch.tensor([near, far], dtype=dtype)  # model/rendering.py:49 (fp32 tensor of the range)
    z = nf[0] * (1.0 - u) + nf[1] * u
    z = z.reshape(1, -1).repeat(n_rays, 1)
    if noise is not None:
        mid = 0.5 * (z[:, 1:] + z[:, :-1])
        hi = torch.cat([mid, z[:, -1:]], dim=-1)
        lo = torch.cat([z[:, :1], mid], dim=-1)
        z = lo + (hi - lo) * noise
    return z


def posenc(x, levels):
    """[x, sin(2^0 x), cos(2^0 x), ..., sin(2^(L-1) x), cos(2^(L-1) x)].  model/official_nerf.py:99-119."""
    out = [x]
    for i in range(levels):
        a = (2.0 ** i) * x
        out += [torch.sin(a), torch.cos(a)]
    return torch.cat(out, dim=-1)


def mlp_trunk(P, pts):
    """8x256 trunk with the skip at layer 5.  model/official_nerf.py:60-67 (infer_occ)."""
    enc = posenc(pts, 10)
    h = enc
    for i in (0, 2, 4, 6):
        h = F.relu(F.linear(h, P["layers0.%d.weight" % i], P["layers0.%d.bias" % i]))
    h = torch.cat([h, enc], dim=-1)
    for i in (0, 2, 4, 6):
        h = F.relu(F.linear(h, P["layers1.%d.weight" % i], P["layers1.%d.bias" % i]))
    raw = F.linear(h, P["fc_density.weight"], P["fc_density.bias"])
    return h, raw


def mlp_forward(P, pts, viewdir, occ_activation="softplus", dist_alpha=False):
    """model/official_nerf.py:69-96.  pts, viewdir (M,3) -> rgb (M,3), alpha-or-sigma (M,1)."""
    h, raw = mlp_trunk(P, pts)
    dens = F.softplus(raw) if occ_activation == "softplus" else raw.relu()
    if not dist_alpha:
        dens = 1 - torch.exp(-1.0 * dens)
    feat = F.linear(h, P["fc_feature.weight"], P["fc_feature.bias"])
    x = torch.cat([feat, posenc(viewdir, 4)], dim=-1)
    x = F.relu(F.linear(x, P["rgb_layers.0.weight"], P["rgb_layers.0.bias"]))
    rgb = torch.sigmoid(F.linear(x, P["fc_rgb.weight"], P["fc_rgb.bias"]))
    return rgb, dens


# ----------------------------------------------------------------------------------------------
# Stage 3: alpha compositing
# ----------------------------------------------------------------------------------------------


def composite(alpha, rgb, z, dist_alpha=False, white_background=False):
    """model/rendering.py:113-126,139-141.  alpha (R,S) [sigma if dist_alpha], rgb (R,S,3), z (R,S).
    Returns rgb (R,3), dist (R,), weights (R,S), alpha (R,S)."""
    if dist_alpha:
        delta = torch.cat([z[:, 1:] - z[:, :-1], torch.full_like(z[:, :1], 1e10)], dim=-1)
        alpha = 1 - torch.exp(-1.0 * alpha * delta)
        alpha = torch.cat([alpha[:, :-1], torch.ones_like(alpha[:, :1])], dim=-1)
    ones = torch.ones_like(alpha[:, :1])
    T = torch.cumprod(torch.cat([ones, 1.0 - alpha + EPS_T], dim=-1), dim=-1)[:, :-1]
    w = alpha * T
    out_rgb = (w.unsqueeze(-1) * rgb).sum(dim=-2)
    out_dist = (w * z).sum(dim=-1)
    if white_background:
        out_rgb = out_rgb + (1.0 - w.sum(-1, keepdim=True))
    return out_rgb, out_dist, w, alpha


def ndc_rays(fxfy, near, rays_o, rays_d):
    """get_ndc_rays_fxfy, model/common.py:632-675 (same op order).  rays_o, rays_d (...,3) -> NDC origin, direction."""
    t = -(near + rays_o[..., 2]) / rays_d[..., 2]
    rays_o = rays_o + t[..., None] * rays_d
    ox_oz = rays_o[..., 0] / rays_o[..., 2]
    oy_oz = rays_o[..., 1] / rays_o[..., 2]
    o0 = -1. / (1 / fxfy[0]) * ox_oz
    o1 = -1. / (1 / fxfy[1]) * oy_oz
    o2 = 1. + 2. * near / rays_o[..., 2]
    d0 = -1. / (1 / fxfy[0]) * (rays_d[..., 0] / rays_d[..., 2] - ox_oz)
    d1 = -1. / (1 / fxfy[1]) * (rays_d[..., 1] / rays_d[..., 2] - oy_oz)
    d2 = 1 - o2
    return torch.stack([o0, o1, o2], -1), torch.stack([d0, d1, d2], -1)


def render(P, pixels, depth, camera_mat, world_mat, scale_mat, cfg_r, noise=None, eval_=False,
           occ_activation="softplus", chunk=64000):
    """The `nope_nerf` volume renderer: `uniform` sampling branch (default) and `ndc` (model/rendering.py:92-93,169-181).
    model/rendering.py:36-168.  `cfg_r` is the `rendering` config section.
    pixels (1,R,2), depth (1,R,1).  `noise` replaces the reference's internal torch.rand draw
    (model/rendering.py:190) so both sides see the same jitter; None = add_noise False.
    Returns the reference's out_dict plus the dense (unmasked) tensors and the mask."""
    S = cfg_r["num_points"] - cfg_r.get("outside_steps", 0)
    near, far = cfg_r["depth_range"]
    dist_alpha = cfg_r["dist_alpha"]
    o, d, nrm, d_gt, mask = ray_setup(pixels, depth, camera_mat, world_mat, scale_mat,
                                      normalise_ray=cfg_r["normalise_ray"])
    R = d.shape[0]
    ndc = cfg_r.get("sample_option", "uniform") == "ndc"
    if ndc:
        # model/rendering.py:169-181: focal from camera_mat, near plane 1.0, samples on [0,1], no jitter
        focal = torch.cat([camera_mat[:, 0, 0], camera_mat[:, 1, 1]])
        o_s, d_s = ndc_rays(focal, 1.0, o.reshape(1, 3).expand(R, 3), d)
        z = stratified_z(R, S, 0.0, 1.0, None, dtype=pixels.dtype)
        pts = (o_s.unsqueeze(1) + d_s.unsqueeze(1) * z.unsqueeze(-1)).reshape(-1, 3)
    else:
        z = stratified_z(R, S, near, far, noise, dtype=pixels.dtype)
        pts = (o.reshape(1, 1, 3) + d.unsqueeze(1) * z.unsqueeze(-1)).reshape(-1, 3)
    view = (-d).unsqueeze(1).expand(R, S, 3).reshape(-1, 3)
    if not cfg_r["use_ray_dir"]:
        view = torch.ones_like(view)
    rgbs, alphas = [], []
    for i in range(0, pts.shape[0], chunk):  # model/rendering.py:102-111
        c, a = mlp_forward(P, pts[i:i + chunk], view[i:i + chunk], occ_activation, dist_alpha)
        rgbs.append(c)
        alphas.append(a)
    rgb_s = torch.cat(rgbs).reshape(R, S, 3)
    alpha = torch.cat(alphas).reshape(R, S)
    rgb, dist, w, alpha = composite(alpha, rgb_s, z, dist_alpha, cfg_r["white_background"])
    if eval_ and cfg_r["normalise_ray"]:  # model/rendering.py:144-148
        dist = dist / nrm
        d_gt = d_gt / nrm
    if not eval_:
        depth_pred, depth_gt = dist[mask], d_gt[mask]
    else:
        depth_pred, depth_gt = dist, d_gt
    if ndc:  # model/rendering.py:158-159
        depth_gt = 1 - 1 / depth_gt
        d_gt = 1 - 1 / d_gt
    return {
        "rgb": rgb.reshape(1, -1, 3), "z_vals": z, "normal": None,
        "depth_pred": depth_pred, "depth_gt": depth_gt, "alpha": alpha,
        # extras (not in the reference dict) used by the stage-level parity tests
        "weights": w, "dist_dense": dist, "d_gt_dense": d_gt, "mask": mask,
        "origin": o, "dirs": d, "dir_norm": nrm,
    }


# ----------------------------------------------------------------------------------------------
# Stage 4: rgb / depth / reprojection losses
# ----------------------------------------------------------------------------------------------


def rgb_loss(pred, gt, kind="l2"):
    """sum|.|/R or sum(.)^2/R with R = pred.shape[1].  model/losses.py:28-33."""
    diff = pred - gt
    s = diff.abs().sum() if kind == "l1" else (diff * diff).sum()
    return s / float(pred.shape[1])


def depth_l1(pred, gt):
    """model/losses.py:60-63."""
    return (pred - gt).abs().sum() / float(pred.shape[0])


def _shift_scale(x):
    t = torch.median(x)  # lower median for even counts; gradient to that single element
    s = torch.mean(torch.abs(x - t))
    return (x - t) / s


def depth_invariant(pred, gt):
    """Scale/shift-invariant (DPT-style) depth loss.  model/losses.py:35-58."""
    return F.mse_loss(_shift_scale(pred), _shift_scale(gt))


def masked_mean(diff, mask):
    """model/losses.py:79-87 (mean_on_mask).  mask broadcast over the last dim."""
    m = mask.expand_as(diff)
    if m.sum() > 0:
        return diff[m].sum() / m.sum()
    return torch.tensor(0.0, dtype=diff.dtype)


def photometric_loss(rgb1, rgb2, valid):
    """model/losses.py:152-159 (get_rgb_s_loss), with_ssim False."""
    return masked_mean((rgb1 - rgb2).abs().clamp(0, 1), valid)


def sample_bilinear(img, p):
    """grid_sample(bilinear, zeros padding, align_corners=True).  model/common.py:75-109 as it is
    called at model/training.py:367,376.  img (1,C,H,W), p (1,N,2) in [-1,1] -> (1,N,C)."""
    v = F.grid_sample(img, p.unsqueeze(1), mode="bilinear", align_corners=True)
    return v.squeeze(2).permute(0, 2, 1)


def project(points, camera_mat):
    """model/common.py:436-457 (project_to_cam).  points (1,N,3) -> xy (1,N,2), valid (1,N,1) bool."""
    hom = torch.cat([points, torch.ones_like(points[..., :1])], dim=-1)
    q = (camera_mat @ hom.transpose(1, 2))[:, :3].transpose(1, 2)
    xy = q[..., :2] / q[..., 2:]
    valid = xy.abs().max(dim=-1)[0] <= 1
    return xy, valid.unsqueeze(-1)


# ----------------------------------------------------------------------------------------------
# Stage 5: Chamfer point-cloud loss
# ----------------------------------------------------------------------------------------------


def nearest_idx(src, dst, slice_len=1024):
    """argmin_j ||src_i - dst_j||_2 (first index on ties).  model/losses.py:129-144, evaluated on
    source slices (the function's own split) to bound memory.
    src (3,S), dst (3,D) -> (S,) int64."""
    out = []
    for a in range(0, src.shape[1], slice_len):
        diff = src[:, a:a + slice_len, None] - dst[:, None, :]
        out.append(torch.argmin(torch.linalg.norm(diff, dim=0), dim=1))
    return torch.cat(out)


def point_point_error(X, Y, idx=None):
    """mean_i ||X_i - Y[idx_i]||.  model/losses.py:145-150.  X (3,S), Y (3,D)."""
    if idx is None:
        idx = nearest_idx(X, Y)
    return torch.linalg.norm(X - Y[:, idx], dim=0).mean()


def chamfer(X, Y):
    """Symmetric 'dense' Chamfer loss.  model/losses.py:116-123.  X, Y (1,P,3)."""
    a, b = X[0].permute(1, 0), Y[0].permute(1, 0)
    return point_point_error(a, b) + point_point_error(b, a)


LOSS_KEYS = ("loss", "loss_rgb", "loss_depth", "l2_mean", "loss_dist_1st", "loss_dist_2nd",
             "loss_pc", "loss_rgb_s", "loss_depth_consistency", "loss_t_cycle")


def total_loss(weights, rgb_pred, rgb_gt, depth_pred=None, depth_gt=None, X=None, Y=None,
               rgb_pc1=None, rgb_pc1_proj=None, valid_points=None, rgb_loss_type="l2",
               depth_loss_type="l1"):
    """Weighted sum with zero-weight terms skipped.  model/losses.py:164-228 (the terms whose weight
    is [0,0] in every shipped config -- pose smoothness, depth consistency, t-cycle -- are 0 here)."""
    zero = torch.tensor(0.0, dtype=rgb_gt.dtype)
    l_rgb = rgb_loss(rgb_pred, rgb_gt, rgb_loss_type) if weights["rgb_weight"] != 0.0 else zero
    if weights["depth_weight"] != 0.0:
        l_d = depth_l1(depth_pred, depth_gt) if depth_loss_type == "l1" else depth_invariant(depth_pred, depth_gt)
    else:
        l_d = zero
    l_pc = chamfer(X, Y) if weights["pc_weight"] != 0.0 else zero
    l_s = photometric_loss(rgb_pc1, rgb_pc1_proj, valid_points) if weights["rgb_s_weight"] != 0.0 else zero
    l2m = F.mse_loss(rgb_pred, rgb_gt) if (weights["rgb_weight"] != 0.0 or weights["depth_weight"] != 0.0) else zero
    loss = (weights["rgb_weight"] * l_rgb + weights["depth_weight"] * l_d
            + weights["pc_weight"] * l_pc + weights["rgb_s_weight"] * l_s)
    return {"loss": loss, "loss_rgb": l_rgb, "loss_depth": l_d, "l2_mean": l2m, "loss_dist_1st": zero,
            "loss_dist_2nd": zero, "loss_pc": l_pc, "loss_rgb_s": l_s,
            "loss_depth_consistency": zero, "loss_t_cycle": zero}


# ----------------------------------------------------------------------------------------------
# The training step's loss (the caller of all five stages)
# ----------------------------------------------------------------------------------------------


def anneal(start, end, anneal_start, anneal_epochs, current):
    """model/training.py:204-212."""
    if current <= anneal_start:
        return start
    if current >= anneal_start + anneal_epochs:
        return end
    return start + (end - start) * (current - anneal_start) / anneal_epochs


def step_weights(cfg_t, epoch, scheduling_start):
    """model/training.py:225-232."""
    names = ["rgb_weight", "depth_weight", "pc_weight", "rgb_s_weight", "depth_consistency_weight",
             "weight_dist_2nd_loss", "weight_dist_1st_loss", "t_cycle_weight"]
    w = {k: anneal(cfg_t[k][0], cfg_t[k][1], scheduling_start, cfg_t["annealing_epochs"], epoch) for k in names}
    kind = "l1" if epoch < cfg_t["annealing_epochs"] + scheduling_start else "l2"
    return w, kind


def backproject_cloud(depth_img, camera_mat, res, nl):
    """Nearest-resize the depth map to `res`, clamp below `nl` (in place in the reference: gradient is
    cut there), unproject with identity world/scale.  model/training.py:354-362."""
    _, p = pixel_grid(res[0], res[1], dtype=depth_img.dtype)
    d = F.interpolate(depth_img, res, mode="nearest")
    d = torch.where(d < nl, torch.full_like(d, nl), d)
    return unproject(p, d.reshape(1, -1, 1), camera_mat), p


def train_loss(batch, P, pose, dist, cfg, ray_idx, noise, epoch=0, scheduling_start=None):
    """Restatement of Trainer.compute_loss, model/training.py:214-416, for the default branches
    (learned pose + distortion, no learned focal, detach_ref_img True; shift_first either way).
    `pose` = dict(r, t, init_c2w|None); `dist` = dict(global_scales, global_shifts).
    `ray_idx` / `noise` replace the reference's randperm / rand draws."""
    ct, cr = cfg["training"], cfg["rendering"]
    if scheduling_start is None:
        scheduling_start = ct["scheduling_start"]
    weights, rgb_kind = step_weights(ct, epoch, scheduling_start)
    render_model = weights["rgb_weight"] != 0.0 or weights["depth_weight"] != 0.0
    use_ref = weights["pc_weight"] != 0.0 or weights["rgb_s_weight"] != 0.0 or weights["t_cycle_weight"] != 0.0
    nl = ct["nearest_limit"]
    img = batch["img"]
    depth = (batch.get("img.dpt") if batch.get("img.dpt") is not None else batch["img.depth"]).unsqueeze(1)
    K, Smat = batch["img.camera_mat"], batch["img.scale_mat"]
    idx = int(batch["img.idx"])
    n_cams = pose["r"].shape[0]
    _, _, h, w = img.shape

    def c2w_of(i):
        init = None if pose.get("init_c2w") is None else pose["init_c2w"][i]
        return pose_c2w(pose["r"][i], pose["t"][i], init)

    c2w = c2w_of(idx)
    world_mat = torch.inverse(c2w).unsqueeze(0)
    scale_in, shift_in = distortion(dist["global_scales"], dist["global_shifts"], idx, cfg["distortion"]["fix_scaleN"])
    shift_first = cfg["training"].get("shift_first", False)
    depth = (depth + shift_in) * scale_in if shift_first else depth * scale_in + shift_in  # model/training.py:261-264

    rgb_gt = img.reshape(1, 3, h * w).permute(0, 2, 1)[:, ray_idx]
    _, p_full = pixel_grid(h, w, dtype=img.dtype)
    p = p_full[:, ray_idx]
    kw = {}
    out = None
    if render_model:
        d_sel = depth.reshape(1, 1, -1).permute(0, 2, 1)[:, ray_idx]  # model/network.py:21-24
        out = render(P, p, d_sel, K, world_mat, Smat, cr, noise=None if noise is None else noise[0],
                     occ_activation=cfg["model"]["occ_activation"], chunk=cr["n_max_network_queries"])
    if use_ref:
        ridx = int(batch["img.ref_idxs"])
        dref = (batch.get("img.ref_dpts") if batch.get("img.ref_dpts") is not None
                else batch["img.ref_depths"]).unsqueeze(1)
        c2w_ref = c2w_of(ridx)
        scale_ref, shift_ref = distortion(dist["global_scales"], dist["global_shifts"], ridx,
                                          cfg["distortion"]["fix_scaleN"])
        dref = scale_ref * (dref + shift_ref) if shift_first else scale_ref * dref + shift_ref  # model/training.py:312-315
        if ct["detach_ref_img"]:
            c2w_ref, scale_ref, shift_ref, dref = c2w_ref.detach(), scale_ref.detach(), shift_ref.detach(), dref.detach()
        ref_Rt = torch.inverse(c2w_ref).unsqueeze(0)
        if idx < n_cams - 1:  # model/training.py:323-352
            d1, d2, img1, img2 = depth, dref, img, batch["img.ref_imgs"]
            rel = ref_Rt @ torch.inverse(world_mat)
            scale1 = scale_in
        else:
            d1, d2, img1, img2 = dref, depth, batch["img.ref_imgs"], img
            rel = world_mat @ torch.inverse(ref_Rt)
            scale1 = scale_ref
        R_rel, t_rel = rel[:, :3, :3], rel[:, :3, 3]
        res = (int(depth.shape[2] / ct["pc_ratio"]), int(depth.shape[3] / ct["pc_ratio"]))
        pc1, p_pc = backproject_cloud(d1, K, res, nl)
        pc2, _ = backproject_cloud(d2, K, res, nl)
        if weights["rgb_s_weight"] != 0.0:  # model/training.py:364-382
            im1 = F.interpolate(img1, res, mode="bilinear")
            im2 = F.interpolate(img2, res, mode="bilinear")
            rgb_pc1 = sample_bilinear(im1, p_pc)
            rot = (pc1.detach().clone() if ct["detach_rgbs_scale"] else pc1) @ R_rel.transpose(1, 2) + t_rel
            behind = (-rot[:, :, 2:] < nl).expand_as(rot)
            rot = torch.where(behind, torch.full_like(rot, nl), rot)
            p_rep, valid = project(rot, K)
            rgb_proj = sample_bilinear(im2, p_rep)
            kw.update(rgb_pc1=rgb_pc1.reshape(1, res[0], res[1], 3), rgb_pc1_proj=rgb_proj.reshape(1, res[0], res[1], 3),
                      valid_points=valid.reshape(1, res[0], res[1], 1))
        if ct["scale_pcs"]:
            pc1, pc2 = pc1 / scale1, pc2 / scale1
        kw.update(X=pc1 @ R_rel.transpose(1, 2) + t_rel, Y=pc2)
    d_gt = out["depth_gt"] if out is not None else None
    if out is not None and ct["detach_gt_depth"]:
        d_gt = d_gt.detach()
    ld = total_loss(weights, out["rgb"] if out is not None else None, rgb_gt,
                    out["depth_pred"] if out is not None else None, d_gt,
                    rgb_loss_type=rgb_kind, depth_loss_type=ct["depth_loss_type"], **kw)
    ld["scale"], ld["shift"] = scale_in, shift_in
    return ld, out, kw


# ----------------------------------------------------------------------------------------------
# Geometry path: density gradient, sphere tracing, phong shading
# ----------------------------------------------------------------------------------------------


def occupancy(P, pts, occ_activation="softplus"):
    """OfficialStaticNerf.forward(only_occupancy=True), model/official_nerf.py:69-85 (dist_alpha False)."""
    _, raw = mlp_trunk(P, pts)
    dens = F.softplus(raw) if occ_activation == "softplus" else raw.relu()
    return 1 - torch.exp(-1.0 * dens)


def density_gradient(P, pts):
    """OfficialStaticNerf.gradient, model/official_nerf.py:46-58: -(d raw_density / d p), (M,1,3)."""
    with torch.enable_grad():
        p = pts.detach().clone().requires_grad_(True)
        _, y = mlp_trunk(P, p)
        g = torch.autograd.grad(y, p, torch.ones_like(y))[0]
    return -g.detach().unsqueeze(1)


def sphere_intersection(cam_loc, ray_directions, r=1.0):
    """get_sphere_intersection, model/rendering.py:440-460.  cam_loc (3,), ray_directions (R,3) -> (R,2), mask (R,)."""
    dot = ray_directions @ cam_loc
    under = dot ** 2 - (cam_loc.norm(2) ** 2 - r ** 2)
    mask = under > 0
    out = torch.zeros(ray_directions.shape[0], 2)
    out[mask] = torch.sqrt(under[mask]).unsqueeze(-1) * torch.tensor([-1.0, 1.0])
    out[mask] -= dot[mask].unsqueeze(-1)
    return out.clamp_min(0.0), mask


def ray_marching(P, cam, dirs, tau=0.5, n_steps=512, n_secant_steps=8, depth_range=(0.0, 2.4), rad=1.0):
    """Renderer.ray_marching + secant, model/rendering.py:274-419.  cam (3,), dirs (R,3) -> d (R,)."""
    R = dirs.shape[0]
    with torch.no_grad():
        d_far = sphere_intersection(cam, dirs, rad)[0][:, 1]
        u = torch.linspace(0, 1, steps=n_steps).view(1, n_steps)
        d_prop = depth_range[0] * (1.0 - u) + d_far.view(R, 1) * u
        p = cam.view(1, 1, 3) + dirs.unsqueeze(1) * d_prop.unsqueeze(-1)
        val = (occupancy(P, p.reshape(-1, 3)).reshape(R, n_steps) - tau)
        mask0 = val[:, 0] < 0
        sign = torch.cat([torch.sign(val[:, :-1] * val[:, 1:]), torch.ones(R, 1)], dim=-1)
        cost = sign * torch.arange(n_steps, 0, -1).float()
        values, indices = torch.min(cost, -1)
        ar = torch.arange(R)
        mask = (values < 0) & (val[ar, indices] < 0) & mask0
        hi = torch.clamp(indices + 1, max=n_steps - 1)
        d_low, f_low = d_prop[ar, indices][mask], val[ar, indices][mask]
        d_high, f_high = d_prop[ar, hi][mask], val[ar, hi][mask]
        out = torch.ones(R)
        if d_low.shape[0]:
            dm = dirs[mask]
            d_pred = -f_low * (d_high - d_low) / (f_high - f_low) + d_low
            for _ in range(n_secant_steps):  # model/rendering.py:401-418
                f_mid = occupancy(P, cam.view(1, 3) + d_pred.unsqueeze(-1) * dm).reshape(-1) - tau
                low = f_mid < 0
                d_low, f_low = torch.where(low, d_pred, d_low), torch.where(low, f_mid, f_low)
                d_high, f_high = torch.where(low, d_high, d_pred), torch.where(low, f_high, f_mid)
                d_pred = -f_low * (d_high - d_low) / (f_high - f_low) + d_low
            out[mask] = d_pred
        out[~mask] = float("inf")
        out[~mask0] = 0
    return out


def phong(P, cam, dirs, rad):
    """Renderer.phong_renderer, model/rendering.py:203-271 -> rgb (R,3) shaded, rgb_surf (R,3), d (R,)."""
    d = ray_marching(P, cam, dirs, rad=rad)
    with torch.no_grad():
        zero_occ, pred = d == 0, torch.isfinite(d)
        dists = torch.ones_like(d)
        dists[pred] = d[pred]
        dists[zero_occ] = 0.0
        obj = pred & ~zero_occ
        pts = cam.view(1, 3) + dirs * dists.unsqueeze(-1)
        rgb, rgb_surf = torch.ones_like(pts), torch.zeros_like(pts)
        if obj.any():
            g = density_gradient(P, pts[obj])[:, 0, :]
            n = g / g.norm(2, 1, keepdim=True)
            light = (cam / cam.norm(2)).view(3, 1)
            diffuse = torch.mm(n, light).clamp_min(0).repeat(1, 3) * 0.7
            rgb[obj] = (0.3 + diffuse).clamp_max(1.0)
            rgb_surf[obj] = mlp_forward(P, pts[obj], -dirs[obj])[0]
    return rgb, rgb_surf, d
