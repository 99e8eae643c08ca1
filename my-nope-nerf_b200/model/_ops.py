"""torch.autograd.Function wrappers: one per C-ABI stage, forward and hand-written backward both
call into libnope_nerf_b200.so on torch's current CUDA stream.  No torch math on the data path."""
import ctypes as C

import torch

from . import _cabi as K
from ._cabi import call, f32, ptr, stream


def _new(ref, *shape, dtype=torch.float32):
    return torch.empty(shape, dtype=dtype, device=ref.device)


# ------------------------------------------------------------------------------------------------
# Stage 1
# ------------------------------------------------------------------------------------------------
class PoseC2W(torch.autograd.Function):
    """LearnPose.forward (model/poses.py:23-31)."""

    @staticmethod
    def forward(ctx, r, t, init_c2w, cam_id):
        r, t = f32(r), f32(t)
        init = f32(init_c2w) if init_c2w is not None else None
        out = _new(r, 4, 4)
        call("nnb_pose_fwd", ptr(r), ptr(t), ptr(init), int(cam_id), ptr(out), stream())
        ctx.save_for_backward(r, t, init if init is not None else torch.empty(0, device=r.device))
        ctx.cam_id, ctx.has_init = int(cam_id), init is not None
        return out

    @staticmethod
    def backward(ctx, g):
        r, t, init = ctx.saved_tensors
        dr, dt = torch.zeros_like(r), torch.zeros_like(t)
        rows = _new(r, 6)
        call("nnb_pose_bwd", ptr(r), ptr(t), ptr(init) if ctx.has_init else None, ctx.cam_id, ptr(f32(g)),
             C.c_void_p(rows.data_ptr()), C.c_void_p(rows.data_ptr() + 12), stream())
        dr[ctx.cam_id] = rows[:3]
        dt[ctx.cam_id] = rows[3:]
        return dr, dt, None, None


class Mat4Inverse(torch.autograd.Function):
    """torch.inverse on one 4x4 (model/training.py:257 and friends)."""

    @staticmethod
    def forward(ctx, m):
        shape = m.shape
        m = f32(m).reshape(4, 4)
        out = torch.empty_like(m)
        call("nnb_mat4_inv_fwd", ptr(m), ptr(out), stream())
        ctx.save_for_backward(out)
        ctx.shape = shape
        return out.reshape(shape)

    @staticmethod
    def backward(ctx, g):
        (inv,) = ctx.saved_tensors
        dm = torch.empty_like(inv)
        call("nnb_mat4_inv_bwd", ptr(inv), ptr(f32(g).reshape(4, 4)), ptr(dm), stream())
        return dm.reshape(ctx.shape)


class Distortion(torch.autograd.Function):
    """Learn_Distortion.forward (model/distortions.py:19-27) -> tensor [scale, shift]."""

    @staticmethod
    def forward(ctx, scales, shifts, cam_id, fix_scaleN):
        scales, shifts = f32(scales), f32(shifts)
        out = _new(scales, 2)
        n = scales.shape[0]
        call("nnb_distortion_fwd", ptr(scales), ptr(shifts), int(cam_id), n, int(bool(fix_scaleN)), ptr(out), stream())
        ctx.save_for_backward(scales)
        ctx.args = (int(cam_id), n, int(bool(fix_scaleN)))
        ctx.set_materialize_grads(False)
        return out

    @staticmethod
    def backward(ctx, g):
        """WHICH parameters receive a gradient follows the reference (torch.optim.Adam skips a parameter whose .grad is None and
        keeps moving one whose gradient is zero): the pinned scale of the last camera is a constant there (model/distortions.py:
        23-24), so `global_scales` gets NO gradient from this call, not a zero one.  (The 0.01 clamp of :21-22 is decided on the
        device here, without the reference's host read: a clamped scale yields a zero row instead of no gradient.)"""
        if g is None:
            return None, None, None, None
        (scales,) = ctx.saved_tensors
        cam, n, fix = ctx.args
        pinned = bool(fix) and cam == n - 1
        ds, dsh = (None if pinned else torch.zeros_like(scales)), torch.zeros_like(scales)
        row = _new(scales, 2)
        call("nnb_distortion_bwd", ptr(scales), cam, n, fix, ptr(f32(g)), C.c_void_p(row.data_ptr()),
             C.c_void_p(row.data_ptr() + 4), stream())
        if ds is not None:
            ds[cam, 0] = row[0]
        dsh[cam, 0] = row[1]
        return ds, dsh, None, None


def pixel_grid(h, w, device):
    loc = torch.empty(h * w, 2, dtype=torch.int64, device=device)
    scaled = torch.empty(h * w, 2, dtype=torch.float32, device=device)
    call("nnb_pixel_grid", h, w, ptr(loc), ptr(scaled), stream())
    return loc, scaled


class Mat4Mul(torch.autograd.Function):
    """a @ b on 4x4 matrices (the relative pose of model/training.py:329,344) without a library GEMM."""

    @staticmethod
    def forward(ctx, a, b):
        a, b = f32(a).reshape(4, 4), f32(b).reshape(4, 4)
        out = torch.empty_like(a)
        call("nnb_mat4_mul_fwd", ptr(a), ptr(b), ptr(out), stream())
        ctx.save_for_backward(a, b)
        return out

    @staticmethod
    def backward(ctx, g):
        a, b = ctx.saved_tensors
        da = torch.empty_like(a) if ctx.needs_input_grad[0] else None
        db = torch.empty_like(b) if ctx.needs_input_grad[1] else None
        if da is None and db is None:
            return None, None
        call("nnb_mat4_mul_bwd", ptr(a), ptr(b), ptr(f32(g).reshape(4, 4)), ptr(da), ptr(db), stream())
        return da, db


def ray_choice(n_total, n_pick, device):
    """n_pick distinct pixel indices of [0, n_total), uniformly scattered: the draw of model/training.py:277
    (`torch.randperm(h*w)[:n_points]`) without sorting h*w keys.  The six round keys come from torch's generator on `device`
    (graph-safe), so `torch.manual_seed` fixes the choice."""
    keys = torch.randint(0, 2 ** 31 - 1, (6,), device=device, dtype=torch.int64)
    out = torch.empty(int(n_pick), device=device, dtype=torch.int64)
    call("nnb_ray_choice", ptr(keys), int(n_total), int(n_pick), ptr(out), stream())
    return out


def copy_many(pairs, scalars=()):
    """One launch for a list of (dst, src) device tensors (same nbytes each pair, contiguous) and up to two
    (dst int64 device tensor, python int) scalars: the batch staging of Trainer.train_step_graphed."""
    for i in range(0, max(len(pairs), 1), K.COPY_MAX):
        part = pairs[i:i + K.COPY_MAX]
        n = len(part)
        src = (C.c_void_p * max(n, 1))(*[s.data_ptr() for _, s in part])
        dst = (C.c_void_p * max(n, 1))(*[d.data_ptr() for d, _ in part])
        nb = (C.c_ulonglong * max(n, 1))(*[d.numel() * d.element_size() for d, _ in part])
        sc = scalars if i == 0 else ()
        sd = (C.c_void_p * 2)(*([t.data_ptr() for t, _ in sc] + [None] * (2 - len(sc))))
        sv = (C.c_int64 * 2)(*([int(v) for _, v in sc] + [0] * (2 - len(sc))))
        call("nnb_copy_many", src, dst, nb, n, sd, sv, stream())


class GatherRays(torch.autograd.Function):
    """pixels / rgb_gt / depth*scale+shift at ray_idx (model/training.py:264,285-288; model/network.py:21-24).
    scale_shift: tensor [scale, shift] on device or None."""

    @staticmethod
    def forward(ctx, ray_idx, img, depth_img, scale_shift, h, w):
        ray_idx = ray_idx.contiguous()
        n = ray_idx.numel()
        img = f32(img) if img is not None else None
        depth_img = f32(depth_img) if depth_img is not None else None
        ss = f32(scale_shift) if scale_shift is not None else None
        pixels = _new(ray_idx, n, 2)
        rgb = _new(ray_idx, n, 3) if img is not None else None
        depth = _new(ray_idx, n) if depth_img is not None else None
        call("nnb_gather_rays_fwd", ptr(ray_idx), n, int(h), int(w), ptr(img), ptr(depth_img), ptr(ss), ptr(pixels), ptr(rgb),
             ptr(depth), stream())
        ctx.save_for_backward(ray_idx, depth_img if depth_img is not None else torch.empty(0, device=ray_idx.device))
        ctx.has_ss = ss is not None
        ctx.set_materialize_grads(False)
        ctx.mark_non_differentiable(pixels)
        if rgb is not None:
            ctx.mark_non_differentiable(rgb)
        return pixels, rgb, depth

    @staticmethod
    def backward(ctx, gp, gr, gd):
        ray_idx, depth_img = ctx.saved_tensors
        dss = None
        if ctx.has_ss and gd is not None:
            dss = _new(depth_img, 2)
            call("nnb_gather_rays_bwd", ptr(ray_idx), ray_idx.numel(), ptr(depth_img), ptr(f32(gd)), ptr(dss), stream())
        return None, None, None, dss, None, None


class RayGen(torch.autograd.Function):
    """model/common.py:112-237 + model/rendering.py:53-80."""

    @staticmethod
    def forward(ctx, pixels, depth, camera_mat, world_mat, scale_mat, normalise):
        ctx.depth_shape = depth.shape if depth is not None else None
        pixels = f32(pixels).reshape(-1, 2)
        n = pixels.shape[0]
        depth = f32(depth).reshape(-1) if depth is not None else None
        Km, Wm, Sm = f32(camera_mat).reshape(4, 4), f32(world_mat).reshape(4, 4), f32(scale_mat).reshape(4, 4)
        A, origin = _new(pixels, 4, 4), _new(pixels, 3)
        dirs, nrm, dgt = _new(pixels, n, 3), _new(pixels, n), _new(pixels, n)
        mask = _new(pixels, n, dtype=torch.bool)
        call("nnb_raygen_fwd", ptr(pixels), ptr(depth), ptr(Km), ptr(Wm), ptr(Sm), n, int(bool(normalise)), ptr(A), ptr(origin),
             ptr(dirs), ptr(nrm), ptr(dgt), ptr(mask), stream())
        ctx.save_for_backward(pixels, depth if depth is not None else torch.empty(0, device=pixels.device), Km, Wm, Sm, A)
        ctx.has_depth, ctx.normalise = depth is not None, int(bool(normalise))
        ctx.shapes = (camera_mat.shape, world_mat.shape, scale_mat.shape, None)
        ctx.mark_non_differentiable(mask, nrm)
        ctx.set_materialize_grads(False)   # an unused output arrives as None, not as zeros
        return origin, dirs, nrm, dgt, mask

    @staticmethod
    def backward(ctx, g_origin, g_dirs, g_nrm, g_dgt, g_mask):
        pixels, depth, Km, Wm, Sm, A = ctx.saved_tensors
        n = pixels.shape[0]
        # the depth input reaches only the ground-truth ray distance (model/rendering.py:65-70): when no loss reads that output
        # the depth -- and the distortion parameters behind it -- get NO gradient, as in the reference, rather than zeros
        d_depth = _new(pixels, n) if (ctx.has_depth and g_dgt is not None) else None
        dK, dW, dS = _new(pixels, 4, 4), _new(pixels, 4, 4), _new(pixels, 4, 4)
        scratch = _new(pixels, 16)
        call("nnb_raygen_bwd", ptr(pixels), ptr(depth) if ctx.has_depth else None, ptr(Km), ptr(Wm), ptr(Sm), ptr(A), n, ctx.normalise,
             ptr(f32(g_dirs)) if g_dirs is not None else None, ptr(f32(g_origin)) if g_origin is not None else None,
             ptr(f32(g_dgt)) if g_dgt is not None else None, ptr(d_depth), ptr(dK), ptr(dW), ptr(dS), ptr(scratch), stream())
        sk, sw, ss, _ = ctx.shapes
        if d_depth is not None:
            d_depth = d_depth.reshape(ctx.depth_shape)
        return None, d_depth, dK.reshape(sk), dW.reshape(sw), dS.reshape(ss), None


# ------------------------------------------------------------------------------------------------
# Stages 2+3
# ------------------------------------------------------------------------------------------------
def make_cfg(n_rays, n_samples, near, far, occ_softplus, dist_alpha, white_bkgd, use_dir, precision):
    return K.RenderCfg(int(n_rays), int(n_samples), float(near), float(far), int(occ_softplus), int(dist_alpha),
                       int(white_bkgd), int(use_dir), int(precision))


def workspace_bytes(cfg, save):
    return int(K.lib().nnb_render_workspace_bytes(C.byref(cfg), int(save)))


_ws_cache = {}


def _workspace(device, nbytes, tag):
    """Reused scratch buffer per (device, tag): the library never allocates."""
    key = (str(device), tag)
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=device)
        _ws_cache[key] = buf
    return buf


# Stream for the weight-gradient kernel of the render backward (nnb_ray_extras.wgrad_stream).  None: everything on the current
# stream.  Whoever sets it (Trainer.train_step_graphed) joins it before reading parameter gradients.
wgrad_stream = None
wgrad_forked = False   # set when a backward handed wgrad_stream to the library since the flag was last cleared


def _extras(ray_o, view, d_ray_o=None, d_view=None, backward=False):
    def v(t):
        return ptr(t).value if t is not None else None
    global wgrad_forked
    ws = wgrad_stream.cuda_stream if (backward and wgrad_stream is not None) else None
    wgrad_forked = wgrad_forked or ws is not None
    return K.RayExtras(v(ray_o), v(view), v(d_ray_o), v(d_view), ws)


def _release_ws(ctx, flat):
    """Drop the saved-activation workspace of a render node after its backward was enqueued.  With the weight-gradient kernel
    forked onto `wgrad_stream` that kernel still reads the workspace (and writes `flat`) after the launching stream has moved
    on: tell the caching allocator, so that the blocks are not handed to a later allocation of the launching stream before
    the fork is joined (inside a capture, blocks with a recorded stream use stay reserved until the capture ends)."""
    if wgrad_stream is not None:
        ctx.ws.record_stream(wgrad_stream)
        flat.record_stream(wgrad_stream)
    ctx.ws = None


def render_nograd(cfg, params, packed, origin, dirs, noise, want_aux=True, ray_o=None, view=None):
    """Forward only (eval / render loops).  Returns rgb, dist, weights, alpha, z (aux None if not wanted).
    ray_o / view: per-ray origins and view directions of the NDC branch (nnb_ray_extras)."""
    R, S = cfg.n_rays, cfg.n_samples
    dev = dirs.device
    origin, dirs = f32(origin), f32(dirs)
    ray_o = f32(ray_o) if ray_o is not None else None
    view = f32(view) if view is not None else None
    noise = f32(noise).reshape(R, S) if noise is not None else None
    params = [f32(p) for p in params]
    rgb, dist = torch.empty(R, 3, device=dev), torch.empty(R, device=dev)
    full = want_aux is True                      # want_aux == 'alpha': only the per-sample alpha
    w = torch.empty(R, S, device=dev) if full else None
    a = torch.empty(R, S, device=dev) if want_aux else None
    z = torch.empty(R, S, device=dev) if full else None
    nbytes = workspace_bytes(cfg, 0)
    ws = _workspace(dev, nbytes, "render_eval")
    parr = K.param_ptr_array(params)
    if ray_o is None and view is None:
        call("nnb_render_fwd", C.byref(cfg), parr, ptr(packed), ptr(origin), ptr(dirs), ptr(noise), ptr(rgb), ptr(dist), ptr(w), ptr(a),
             ptr(z), ptr(ws), nbytes, 0, stream())
    else:
        ex = _extras(ray_o, view)
        call("nnb_render_fwd_ex", C.byref(cfg), parr, ptr(packed), ptr(origin), ptr(dirs), C.byref(ex), ptr(noise), ptr(rgb), ptr(dist),
             ptr(w), ptr(a), ptr(z), ptr(ws), nbytes, 0, stream())
    return rgb, dist, w, a, z


class RenderRays(torch.autograd.Function):
    """Stratified sampling + positional encoding + MLP + compositing (model/rendering.py:89-126,
    model/official_nerf.py:60-119).  Inputs after the fixed ones are the 24 parameter tensors."""

    @staticmethod
    def forward(ctx, cfg, packed, origin, dirs, noise, *params):
        R, S = cfg.n_rays, cfg.n_samples
        dev = dirs.device
        origin, dirs = f32(origin), f32(dirs)
        noise = f32(noise).reshape(R, S) if noise is not None else None
        params = [f32(p) for p in params]
        rgb, dist = torch.empty(R, 3, device=dev), torch.empty(R, device=dev)
        w, a, z = torch.empty(R, S, device=dev), torch.empty(R, S, device=dev), torch.empty(R, S, device=dev)
        nbytes = workspace_bytes(cfg, 1)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)  # owned by this graph node until backward
        parr = K.param_ptr_array(params)
        call("nnb_render_fwd", C.byref(cfg), parr, ptr(packed), ptr(origin), ptr(dirs), ptr(noise), ptr(rgb), ptr(dist), ptr(w),
             ptr(a), ptr(z), ptr(ws), nbytes, 1, stream())
        ctx.cfg, ctx.packed, ctx.ws, ctx.nbytes = cfg, packed, ws, nbytes
        ctx.save_for_backward(origin, dirs, w, a, z, *params)
        ctx.mark_non_differentiable(z)
        ctx.set_materialize_grads(False)   # unused outputs (weights, alpha) arrive as None in backward, not as zero fills
        return rgb, dist, w, a, z

    @staticmethod
    def backward(ctx, g_rgb, g_dist, g_w, g_a, g_z):
        origin, dirs, w, a, z, *params = ctx.saved_tensors
        cfg = ctx.cfg
        dev = dirs.device
        flat = torch.zeros(sum(p.numel() for p in params), device=dev)  # one fill instead of 24
        grads, o = [], 0
        for p in params:
            grads.append(flat[o:o + p.numel()].view_as(p))
            o += p.numel()
        d_origin, d_dirs = torch.empty(3, device=dev), torch.empty_like(dirs)
        parr, garr = K.param_ptr_array(params), K.param_ptr_array(grads)

        def opt(g):
            return ptr(f32(g)) if g is not None else None
        ex = _extras(None, None, backward=True)
        call("nnb_render_bwd_ex", C.byref(cfg), parr, ptr(ctx.packed), ptr(origin), ptr(dirs), C.byref(ex), ptr(w), ptr(a), ptr(z),
             opt(g_rgb), opt(g_dist), opt(g_a), opt(g_w), garr, ptr(d_origin), ptr(d_dirs), ptr(ctx.ws), ctx.nbytes, stream())
        _release_ws(ctx, flat)
        return (None, None, d_origin, d_dirs, None) + tuple(grads)


class RenderRaysEx(torch.autograd.Function):
    """RenderRays with per-ray origins and separate view directions (the `sample_ndc` branch, model/rendering.py:169-181):
    points = ray_o[r] + dirs[r] * z, view-direction encoding of -view[r]."""

    @staticmethod
    def forward(ctx, cfg, packed, ray_o, dirs, view, noise, *params):
        R, S = cfg.n_rays, cfg.n_samples
        dev = dirs.device
        ray_o, dirs, view = f32(ray_o), f32(dirs), f32(view)
        noise = f32(noise).reshape(R, S) if noise is not None else None
        params = [f32(p) for p in params]
        rgb, dist = torch.empty(R, 3, device=dev), torch.empty(R, device=dev)
        w, a, z = torch.empty(R, S, device=dev), torch.empty(R, S, device=dev), torch.empty(R, S, device=dev)
        nbytes = workspace_bytes(cfg, 1)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        parr = K.param_ptr_array(params)
        ex = _extras(ray_o, view)
        call("nnb_render_fwd_ex", C.byref(cfg), parr, ptr(packed), ptr(ray_o), ptr(dirs), C.byref(ex), ptr(noise), ptr(rgb), ptr(dist),
             ptr(w), ptr(a), ptr(z), ptr(ws), nbytes, 1, stream())
        ctx.cfg, ctx.packed, ctx.ws, ctx.nbytes = cfg, packed, ws, nbytes
        ctx.save_for_backward(ray_o, dirs, view, w, a, z, *params)
        ctx.mark_non_differentiable(z)
        ctx.set_materialize_grads(False)   # unused outputs (weights, alpha) arrive as None in backward, not as zero fills
        return rgb, dist, w, a, z

    @staticmethod
    def backward(ctx, g_rgb, g_dist, g_w, g_a, g_z):
        ray_o, dirs, view, w, a, z, *params = ctx.saved_tensors
        cfg = ctx.cfg
        dev = dirs.device
        flat = torch.zeros(sum(p.numel() for p in params), device=dev)
        grads, o = [], 0
        for p in params:
            grads.append(flat[o:o + p.numel()].view_as(p))
            o += p.numel()
        d_origin, d_dirs = torch.empty(3, device=dev), torch.empty_like(dirs)
        d_ray_o, d_view = torch.empty_like(ray_o), torch.empty_like(view)
        parr, garr = K.param_ptr_array(params), K.param_ptr_array(grads)
        ex = _extras(ray_o, view, d_ray_o, d_view, backward=True)

        def opt(g):
            return ptr(f32(g)) if g is not None else None
        call("nnb_render_bwd_ex", C.byref(cfg), parr, ptr(ctx.packed), ptr(ray_o), ptr(dirs), C.byref(ex), ptr(w), ptr(a), ptr(z),
             opt(g_rgb), opt(g_dist), opt(g_a), opt(g_w), garr, ptr(d_origin), ptr(d_dirs), ptr(ctx.ws), ctx.nbytes, stream())
        _release_ws(ctx, flat)
        return (None, None, d_ray_o, d_dirs, d_view, None) + tuple(grads)


class NdcRays(torch.autograd.Function):
    """get_ndc_rays_fxfy (model/common.py:632-675) with focal = camera_mat[0,0], camera_mat[1,1] (model/rendering.py:171-173).
    camera_mat (1,4,4) or (4,4), origin (3,), dirs (R,3) -> ndc_o (R,3), ndc_d (R,3)."""

    @staticmethod
    def forward(ctx, camera_mat, origin, dirs, near):
        cm, origin, dirs = f32(camera_mat).reshape(4, 4), f32(origin), f32(dirs)
        R = dirs.shape[0]
        no, nd = torch.empty_like(dirs), torch.empty_like(dirs)
        call("nnb_ndc_rays_fwd", ptr(cm), float(near), ptr(origin), ptr(dirs), R, ptr(no), ptr(nd), stream())
        ctx.save_for_backward(cm, origin, dirs)
        ctx.near, ctx.cm_shape = float(near), camera_mat.shape
        return no, nd

    @staticmethod
    def backward(ctx, g_no, g_nd):
        cm, origin, dirs = ctx.saved_tensors
        R = dirs.shape[0]
        d_rows, d_dirs, d_f = torch.empty_like(dirs), torch.empty_like(dirs), torch.empty(2, device=dirs.device)
        call("nnb_ndc_rays_bwd", ptr(cm), ctx.near, ptr(origin), ptr(dirs), R, ptr(f32(g_no)) if g_no is not None else None,
             ptr(f32(g_nd)) if g_nd is not None else None, ptr(d_rows), ptr(d_dirs), ptr(d_f), stream())
        d_cm = torch.zeros(16, device=dirs.device)
        d_cm[0], d_cm[5] = d_f[0], d_f[1]
        return d_cm.view(ctx.cm_shape), d_rows.sum(0), d_dirs, None


class Composite(torch.autograd.Function):
    """Stand-alone stage 3 (model/rendering.py:116-126)."""

    @staticmethod
    def forward(ctx, alpha, rgb_s, z, dist_alpha, white):
        alpha, rgb_s, z = f32(alpha), f32(rgb_s), f32(z)
        R, S = alpha.shape
        rgb, dist, w, a = _new(alpha, R, 3), _new(alpha, R), _new(alpha, R, S), _new(alpha, R, S)
        call("nnb_composite_fwd", ptr(alpha), ptr(rgb_s), ptr(z), R, S, int(dist_alpha), int(white), ptr(rgb), ptr(dist), ptr(w),
             ptr(a), stream())
        ctx.save_for_backward(alpha, rgb_s, z, w)
        ctx.flags = (int(dist_alpha), int(white))
        return rgb, dist, w

    @staticmethod
    def backward(ctx, g_rgb, g_dist, g_w):
        alpha, rgb_s, z, w = ctx.saved_tensors
        R, S = alpha.shape
        da, dc = torch.empty_like(alpha), torch.empty_like(rgb_s)
        call("nnb_composite_bwd", ptr(alpha), ptr(rgb_s), ptr(z), ptr(w), R, S, ctx.flags[0], ctx.flags[1],
             ptr(f32(g_rgb)) if g_rgb is not None else None, ptr(f32(g_dist)) if g_dist is not None else None,
             ptr(f32(g_w)) if g_w is not None else None, ptr(da), ptr(dc), stream())
        return da, dc, None, None, None


# ------------------------------------------------------------------------------------------------
# Stage 4
# ------------------------------------------------------------------------------------------------
class RgbLoss(torch.autograd.Function):
    """-> tensor [l1_sum/R, l2_sum/R, mse_mean] (model/losses.py:28-33,200)."""

    @staticmethod
    def forward(ctx, pred, gt):
        ctx.shp = pred.shape
        pred, gt = f32(pred).reshape(-1, 3), f32(gt).reshape(-1, 3)
        out = _new(pred, 3)
        call("nnb_rgb_loss_fwd", ptr(pred), ptr(gt), pred.shape[0], ptr(out), stream())
        ctx.save_for_backward(pred, gt)
        return out

    @staticmethod
    def backward(ctx, g):
        pred, gt = ctx.saved_tensors
        d = torch.empty_like(pred)
        call("nnb_rgb_loss_bwd", ptr(pred), ptr(gt), pred.shape[0], ptr(f32(g)), ptr(d), stream())
        return d.reshape(ctx.shp), None


class DepthL1(torch.autograd.Function):
    """Masked mean |pred-gt| on dense tensors (model/losses.py:60-63)."""

    @staticmethod
    def forward(ctx, pred, gt, mask):
        ctx.shp = (pred.shape, gt.shape)
        pred, gt = f32(pred).reshape(-1), f32(gt).reshape(-1)
        mask = mask.contiguous() if mask is not None else None
        out = _new(pred, 2)
        call("nnb_depth_l1_fwd", ptr(pred), ptr(gt), ptr(mask), pred.numel(), ptr(out), stream())
        ctx.save_for_backward(pred, gt, out, mask if mask is not None else torch.empty(0, device=pred.device))
        ctx.has_mask = mask is not None
        return out[0]

    @staticmethod
    def backward(ctx, g):
        pred, gt, out, mask = ctx.saved_tensors
        dp, dg = torch.empty_like(pred), torch.empty_like(gt)
        call("nnb_depth_l1_bwd", ptr(pred), ptr(gt), ptr(mask) if ctx.has_mask else None, pred.numel(), ptr(out),
             ptr(f32(g).reshape(1)), ptr(dp), ptr(dg), stream())
        return dp.reshape(ctx.shp[0]), dg.reshape(ctx.shp[1]), None


def median_select(x, mask=None):
    x = f32(x).reshape(-1)
    val, idx = _new(x, 1), _new(x, 1, dtype=torch.int64)
    call("nnb_median_select", ptr(x), ptr(mask.contiguous()) if mask is not None else None, x.numel(), ptr(val), ptr(idx), stream())
    return val, idx


class DepthInvariant(torch.autograd.Function):
    """Scale/shift-invariant depth loss (model/losses.py:35-58)."""

    @staticmethod
    def forward(ctx, pred, gt, mask):
        ctx.shp = (pred.shape, gt.shape)
        pred, gt = f32(pred).reshape(-1), f32(gt).reshape(-1)
        mask = mask.contiguous() if mask is not None else None
        loss, stats, med = _new(pred, 1), _new(pred, 8), _new(pred, 2, dtype=torch.int64)
        call("nnb_depth_invariant_fwd", ptr(pred), ptr(gt), ptr(mask), pred.numel(), ptr(loss), ptr(stats), ptr(med), stream())
        ctx.save_for_backward(pred, gt, stats, med, mask if mask is not None else torch.empty(0, device=pred.device))
        ctx.has_mask = mask is not None
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        pred, gt, stats, med, mask = ctx.saved_tensors
        dp, dg = torch.empty_like(pred), torch.empty_like(gt)
        call("nnb_depth_invariant_bwd", ptr(pred), ptr(gt), ptr(mask) if ctx.has_mask else None, pred.numel(), ptr(stats), ptr(med),
             ptr(f32(g).reshape(1)), ptr(dp), ptr(dg), stream())
        return dp.reshape(ctx.shp[0]), dg.reshape(ctx.shp[1]), None


class Backproject(torch.autograd.Function):
    """Depth map -> (h4*w4, 3) cloud (model/training.py:354-362)."""

    @staticmethod
    def forward(ctx, depth_img, scale_shift, camera_mat, h4, w4, nl):
        depth_img = f32(depth_img)
        h, w = depth_img.shape[-2:]
        depth_img = depth_img.reshape(h, w)
        ss = f32(scale_shift) if scale_shift is not None else None
        Km = f32(camera_mat).reshape(4, 4)
        pc = _new(depth_img, h4 * w4, 3)
        clamped = _new(depth_img, h4 * w4, dtype=torch.uint8)
        call("nnb_backproject_fwd", ptr(depth_img), h, w, int(h4), int(w4), ptr(ss), ptr(Km), float(nl), ptr(pc), ptr(clamped), stream())
        ctx.save_for_backward(depth_img, Km, clamped)
        ctx.dims, ctx.has_ss = (h, w, int(h4), int(w4)), ss is not None
        return pc

    @staticmethod
    def backward(ctx, g):
        depth_img, Km, clamped = ctx.saved_tensors
        h, w, h4, w4 = ctx.dims
        dss = None
        if ctx.has_ss:
            dss = _new(depth_img, 2)
            call("nnb_backproject_bwd", ptr(depth_img), h, w, h4, w4, ptr(Km), ptr(clamped), ptr(f32(g)), ptr(dss), stream())
        return None, dss, None, None, None, None


def resize_bilinear(img, h4, w4):
    img = f32(img)
    c, h, w = img.shape[-3:]
    out = _new(img, c, h4, w4)
    call("nnb_resize_bilinear", ptr(img.reshape(c, h, w)), c, h, w, int(h4), int(w4), ptr(out), stream())
    return out


class Reproject(torch.autograd.Function):
    """model/training.py:367-382.  Returns rgb_pc1 (P,3), rgb_pc1_proj (P,3), valid (P,) bool."""

    @staticmethod
    def forward(ctx, img1r, img2r, pc1, R_rel, t_rel, camera_mat, nl):
        ctx.pc_shape = pc1.shape
        img1r, img2r, pc1 = f32(img1r), f32(img2r), f32(pc1).reshape(-1, 3)
        Rm, tv, Km = f32(R_rel).reshape(3, 3), f32(t_rel).reshape(3), f32(camera_mat).reshape(4, 4)
        h4, w4 = img2r.shape[-2:]
        P = pc1.shape[0]
        rgb1, rgb2, valid = _new(pc1, P, 3), _new(pc1, P, 3), _new(pc1, P, dtype=torch.bool)
        call("nnb_reproj_fwd", ptr(img1r), ptr(img2r), h4, w4, ptr(pc1), ptr(Rm), ptr(tv), ptr(Km), float(nl), ptr(rgb1), ptr(rgb2),
             ptr(valid), stream())
        ctx.save_for_backward(img2r, pc1, Rm, tv, Km)
        ctx.nl, ctx.rshape, ctx.tshape = float(nl), R_rel.shape, t_rel.shape
        ctx.mark_non_differentiable(rgb1, valid)
        return rgb1, rgb2, valid

    @staticmethod
    def backward(ctx, g1, g2, gv):
        img2r, pc1, Rm, tv, Km = ctx.saved_tensors
        h4, w4 = img2r.shape[-2:]
        d_pc1 = torch.empty_like(pc1)
        acc = _new(pc1, 12)
        call("nnb_reproj_bwd", ptr(img2r), h4, w4, ptr(pc1), ptr(Rm), ptr(tv), ptr(Km), ctx.nl, ptr(f32(g2).reshape(-1, 3)), ptr(d_pc1),
             C.c_void_p(acc.data_ptr()), C.c_void_p(acc.data_ptr() + 36), stream())
        return None, None, d_pc1.reshape(ctx.pc_shape), acc[:9].reshape(ctx.rshape), acc[9:].reshape(ctx.tshape), None, None


class Photometric(torch.autograd.Function):
    """model/losses.py:79-87,152-159."""

    @staticmethod
    def forward(ctx, rgb1, rgb2, valid):
        ctx.shp = (rgb1.shape, rgb2.shape)
        rgb1, rgb2 = f32(rgb1).reshape(-1, 3), f32(rgb2).reshape(-1, 3)
        valid = valid.reshape(-1).contiguous()
        out = _new(rgb1, 2)
        call("nnb_photometric_fwd", ptr(rgb1), ptr(rgb2), ptr(valid), rgb1.shape[0], ptr(out), stream())
        ctx.save_for_backward(rgb1, rgb2, valid, out)
        return out[0]

    @staticmethod
    def backward(ctx, g):
        rgb1, rgb2, valid, out = ctx.saved_tensors
        d1, d2 = torch.empty_like(rgb1), torch.empty_like(rgb2)
        call("nnb_photometric_bwd", ptr(rgb1), ptr(rgb2), ptr(valid), rgb1.shape[0], ptr(out), ptr(f32(g).reshape(1)), ptr(d1), ptr(d2),
             stream())
        return d1.reshape(ctx.shp[0]), d2.reshape(ctx.shp[1]), None


# ------------------------------------------------------------------------------------------------
# Stage 5
# ------------------------------------------------------------------------------------------------
def chamfer_nn(src, dst, want_dist=False):
    """First-argmin nearest neighbour of every src point in dst (model/losses.py:129-144).  (n,3),(m,3) -> (n,) int64."""
    src, dst = f32(src.detach()).reshape(-1, 3), f32(dst.detach()).reshape(-1, 3)
    idx = _new(src, src.shape[0], dtype=torch.int64)
    dist = _new(src, src.shape[0]) if want_dist else None
    nb = int(K.lib().nnb_chamfer_workspace_bytes(src.shape[0], dst.shape[0]))
    ws = _workspace(src.device, nb, "chamfer")
    call("nnb_chamfer_nn", ptr(src), src.shape[0], ptr(dst), dst.shape[0], ptr(idx), ptr(dist), ptr(ws), nb, stream())
    return (idx, dist) if want_dist else idx


class ChamferErr(torch.autograd.Function):
    """mean_i ||X_i - Y[idx_i]|| (model/losses.py:145-150)."""

    @staticmethod
    def forward(ctx, X, Y, idx):
        ctx.shp = (X.shape, Y.shape)
        X, Y = f32(X).reshape(-1, 3), f32(Y).reshape(-1, 3)
        loss = _new(X, 1)
        call("nnb_chamfer_err_fwd", ptr(X), X.shape[0], ptr(Y), ptr(idx), ptr(loss), stream())
        ctx.save_for_backward(X, Y, idx)
        return loss[0]

    @staticmethod
    def backward(ctx, g):
        X, Y, idx = ctx.saved_tensors
        dX, dY = torch.empty_like(X), torch.zeros_like(Y)
        call("nnb_chamfer_err_bwd", ptr(X), X.shape[0], ptr(Y), ptr(idx), ptr(f32(g).reshape(1)), ptr(dX), ptr(dY), stream())
        return dX.reshape(ctx.shp[0]), dY.reshape(ctx.shp[1]), None


class AlignClouds(torch.autograd.Function):
    """X = (pc1 / scale) @ R.T + t, Y = pc2 / scale (model/training.py:394-399) with Rt the 4x4 relative pose; scale may be None."""

    @staticmethod
    def forward(ctx, pc1, pc2, Rt, scale):
        pc1, pc2, Rt = f32(pc1).reshape(-1, 3), f32(pc2).reshape(-1, 3), f32(Rt).reshape(4, 4)
        sc = f32(scale).reshape(1) if scale is not None else None
        X, Y = torch.empty_like(pc1), torch.empty_like(pc2)
        call("nnb_align_clouds_fwd", ptr(pc1), pc1.shape[0], ptr(pc2), pc2.shape[0], ptr(Rt), ptr(sc), ptr(X), ptr(Y), stream())
        ctx.save_for_backward(pc1, pc2, Rt, sc)
        ctx.scale_shape = scale.shape if scale is not None else None
        return X, Y

    @staticmethod
    def backward(ctx, dX, dY):
        pc1, pc2, Rt, sc = ctx.saved_tensors
        d1, d2 = torch.empty_like(pc1), torch.empty_like(pc2)
        dRt = _new(Rt, 4, 4)
        dsc = _new(Rt, 1) if sc is not None else None
        call("nnb_align_clouds_bwd", ptr(pc1), pc1.shape[0], ptr(pc2), pc2.shape[0], ptr(Rt), ptr(sc), ptr(f32(dX).reshape(-1, 3)),
             ptr(f32(dY).reshape(-1, 3)), ptr(d1), ptr(d2), ptr(dRt), ptr(dsc), stream())
        return d1, d2, dRt, (dsc.reshape(ctx.scale_shape) if sc is not None else None)


def adam_step(param, grad, exp_avg, exp_avg_sq, lr, beta1, beta2, eps, step):
    call("nnb_adam_step", ptr(param), ptr(grad), ptr(exp_avg), ptr(exp_avg_sq), param.numel(), float(lr), float(beta1), float(beta2),
         float(eps), int(step), stream())


def adam_step_raw(param_ptr, grad_ptr, m_ptr, v_ptr, n, lr, beta1, beta2, eps, step):
    """Same on raw device addresses (sub-ranges of the flat buffers of model/optim.py)."""
    call("nnb_adam_step", C.c_void_p(param_ptr), C.c_void_p(grad_ptr), C.c_void_p(m_ptr), C.c_void_p(v_ptr), int(n), float(lr),
         float(beta1), float(beta2), float(eps), int(step), stream())


def mlp_points(params, pts, viewdirs, occ_softplus, dist_alpha):
    """OfficialStaticNerf.forward at explicit points (forward only).  -> rgb (M,3) or None, dens (M,1)."""
    pts = f32(pts)
    M = pts.shape[0]
    viewdirs = f32(viewdirs) if viewdirs is not None else None
    rgb = _new(pts, M, 3) if viewdirs is not None else None
    dens = _new(pts, M)
    nbytes = int(K.lib().nnb_mlp_points_workspace_bytes(M))
    ws = _workspace(pts.device, nbytes, "points")
    parr = K.param_ptr_array([f32(p.detach()) for p in params])
    call("nnb_mlp_points_fwd", parr, ptr(pts), ptr(viewdirs), M, int(bool(occ_softplus)), int(bool(dist_alpha)), ptr(rgb), ptr(dens),
         ptr(ws), nbytes, stream())
    return rgb, dens.reshape(M, 1)


def mlp_points_density_grad(params, pts, want_raw=False):
    """-d fc_density(trunk(p)) / d p at explicit points (OfficialStaticNerf.gradient, model/official_nerf.py:46-58).
    pts (M,3) -> grad (M,3) [, raw density (M,)]."""
    pts = f32(pts.detach()).reshape(-1, 3)
    M = pts.shape[0]
    dev = pts.device
    params = [f32(p.detach()) for p in params]
    grad = torch.empty(M, 3, device=dev)
    raw = torch.empty(M, device=dev) if want_raw else None
    nbytes = int(K.lib().nnb_mlp_points_grad_workspace_bytes(M))
    ws = _workspace(dev, nbytes, "mlp_points_grad")
    call("nnb_mlp_points_density_grad", K.param_ptr_array(params), ptr(pts), M, ptr(grad), ptr(raw), ptr(ws), nbytes, stream())
    return (grad, raw) if want_raw else grad


def alloc_packed(device):
    return torch.empty(int(K.lib().nnb_pack_weights_bytes()), dtype=torch.uint8, device=device)


def pack_weights(params, out=None):
    """bf16 tile images of the 24 tensors for the tcgen05 path (refreshed whenever the fp32 masters change); `out`: an
    `alloc_packed` buffer to re-pack in place."""
    packed = out if out is not None else alloc_packed(params[0].device)
    parr = K.param_ptr_array([f32(p.detach()) for p in params])
    call("nnb_pack_weights", parr, ptr(packed), stream())
    return packed
