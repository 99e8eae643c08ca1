"""Renderer -- drop-in for the reference's model/rendering.py (`nope_nerf` technique, `uniform` and `ndc` sampling).

forward(pixels, depth, camera_mat, world_mat, scale_mat, rendering_technique, add_noise=True, eval_=False, it=...)
returns the reference's dict: rgb (1,R,3), z_vals (R,S), normal, depth_pred, depth_gt, alpha (R,S).
Two launches replace the ~60 the reference issues: ray generation, then the fused
sampling + encoding + MLP + compositing render.  Extra keys (`depth_pred_dense`, `depth_gt_dense`, `mask`) let the
loss stack consume the dense tensors + mask without the reference's boolean-index host sync.
"""
import os

import torch
import torch.nn as nn

from . import _cabi, _ops, geometry

# rays per fused-render call when saving for backward (bounds the activation workspace)
MAX_TRAIN_RAYS_FP32 = 8192


class Renderer(nn.Module):
    def __init__(self, model, cfg, device=None, **kwargs):
        super().__init__()
        self._device = device
        self.depth_range = cfg['depth_range']
        self.n_max_network_queries = cfg['n_max_network_queries']  # kept for parity; the fused kernels do not chunk
        self.white_background = cfg['white_background']
        self.cfg = cfg
        self.model = model.to(device)
        # 'bf16' (tcgen05 tensor cores, fp32 accumulate) or 'fp32' (CUDA cores, tight tolerance).  Not a reference option, so it
        # needs no YAML key: keyword > cfg['precision'] > env NNB_PRECISION > 'auto' = bf16 whenever the tensor-core tile fits
        # (one ray = 128 samples), else fp32.  Both are CUDA paths of this library; there is no CPU path behind either.
        prec = kwargs.get('precision', cfg.get('precision', os.environ.get('NNB_PRECISION', 'auto')))
        if prec == 'auto':
            prec = 'bf16' if (cfg['num_points'] - cfg['outside_steps']) == 128 else 'fp32'
        if prec not in ('bf16', 'fp32'):
            raise ValueError("precision must be 'bf16', 'fp32' or 'auto' (got %r)" % (prec,))
        self.precision = prec
        self._packed = None
        self._packed_version = None
        self._pack_event = None

    # -- reference API -------------------------------------------------------------------------
    def forward(self, pixels, depth, camera_mat, world_mat, scale_mat, rendering_technique, add_noise=True,
                eval_=False, it=1000000, noise=None, lazy_mask=False, shard=None):
        if rendering_technique == 'nope_nerf':
            return self.nope_nerf(pixels, depth, camera_mat, world_mat, scale_mat, it=it, add_noise=add_noise,
                                  eval_=eval_, noise=noise, lazy_mask=lazy_mask, shard=shard)
        if rendering_technique == 'phong_renderer':
            return self.phong_renderer(pixels, camera_mat, world_mat, scale_mat, it=it)
        raise ValueError("unknown rendering technique %r" % (rendering_technique,))

    def phong_renderer(self, pixels, camera_mat, world_mat, scale_mat, it):
        """model/rendering.py:203-271: sphere-traced surface, normals from the density gradient, phong shading (geometry.py)."""
        with torch.no_grad():
            ones = torch.ones(1, pixels.shape[1], 1, device=pixels.device)
            origin, dirs, _, _, _ = _ops.RayGen.apply(pixels, ones, camera_mat, world_mat, scale_mat, True)
        return geometry.phong_render(self, origin, dirs, it=it)

    def ray_marching(self, ray0, ray_direction, model, c=None, tau=0.5, n_steps=[128, 129], n_secant_steps=8,
                     depth_range=[0., 2.4], max_points=3500000, rad=1.0):
        """Reference signature (model/rendering.py:274): ray0 (1,R,3) repeated camera centre, ray_direction (1,R,3) -> (1,R)."""
        return geometry.ray_marching(model, ray0[0, 0], ray_direction[0], tau=0.5, n_steps=int(n_steps[0]),
                                     n_secant_steps=n_secant_steps, depth_range=depth_range, max_points=max_points,
                                     rad=rad, renderer=self if model is self.model else None).unsqueeze(0)

    def secant(self, f_low, f_high, d_low, d_high, n_secant_steps, ray0_masked, ray_direction_masked, tau, it=0):
        """Reference signature (model/rendering.py:389-419); the field evaluations are C-ABI kernels (geometry.secant)."""
        return geometry.secant(self.model, f_low, f_high, d_low, d_high, n_secant_steps, ray0_masked, ray_direction_masked, tau)

    def transform_to_homogenous(self, p):
        """(B,N,3) -> (B,N,4): [p, 1] / |p| (model/rendering.py:421-426)."""
        return torch.cat([p, torch.ones_like(p[..., :1])], dim=2) / p.norm(dim=2, keepdim=True)

    # The two sampling routines of the reference (model/rendering.py:169-198) as stand-alone methods.  `nope_nerf` does NOT call
    # them: sampling is the prologue of the fused render kernels (csrc/mlp_tc.cu z_sample, csrc/render.cu); these exist for callers
    # that want the sample positions themselves, and tests/test_gpu_edges.py checks the kernels' z values against them.
    def _sample_points(self, origins, directions, view_dirs, z_val):
        steps = z_val.shape[-1]
        pts = (origins.unsqueeze(-2) + directions.unsqueeze(-2) * z_val.unsqueeze(-1)).reshape(-1, 3)
        view = -view_dirs.unsqueeze(-2).expand(-1, -1, steps, -1).reshape(-1, 3)
        return z_val.reshape(-1, steps, 1), pts, view

    def sample_uniform(self, camera_world, ray_vector, z_val, add_noise, depth_range, noise=None):
        """z_val (B,R,S) in [0,1] -> depths in depth_range, jittered inside their strata when add_noise (`noise`: the uniform
        draws to use instead of torch.rand); returns z (B*R,S,1), points (B*R*S,3), negated view directions (B*R*S,3)."""
        z_val = depth_range[0] * (1. - z_val) + depth_range[1] * z_val
        if add_noise:
            mid = .5 * (z_val[..., 1:] + z_val[..., :-1])
            lower, upper = torch.cat([z_val[..., :1], mid], dim=-1), torch.cat([mid, z_val[..., -1:]], dim=-1)
            if noise is None:
                noise = torch.rand(z_val.shape, device=z_val.device)
            z_val = lower + (upper - lower) * noise
        return self._sample_points(camera_world, ray_vector, ray_vector, z_val)

    def sample_ndc(self, camera_mat, camera_world, ray_vector, z_val, depth_range=[0., 1.]):
        """The LLFF branch: origins and directions warped to NDC (common.get_ndc_rays_fxfy), no jitter."""
        from .common import get_ndc_rays_fxfy
        focal = torch.cat([camera_mat[:, 0, 0], camera_mat[:, 1, 1]])
        o_ndc, d_ndc = get_ndc_rays_fxfy(focal, 1.0, rays_o=camera_world, rays_d=ray_vector)
        z_val = depth_range[0] * (1. - z_val) + depth_range[1] * z_val
        return self._sample_points(o_ndc, d_ndc, ray_vector, z_val)

    def to(self, device):
        new = super().to(device)
        new._device = device
        return new

    # -- implementation --------------------------------------------------------------------------
    def _prec(self):
        return _cabi.PREC_BF16 if self.precision == 'bf16' else _cabi.PREC_FP32

    def _cfg(self, n_rays, n_samples, depth_range=None):
        c = self.cfg
        dr = depth_range if depth_range is not None else c['depth_range']
        return _ops.make_cfg(n_rays, n_samples, dr[0], dr[1],
                             self.model.occ_activation == 'softplus', c['dist_alpha'], c['white_background'],
                             c['use_ray_dir'], self._prec())

    def _packed_weights(self, params):
        """bf16 stage images of the 24 fp32 masters, in ONE persistent buffer that is re-packed in place: its address is
        baked into captured graphs, so it must never be reallocated.  Eagerly the pack runs when a master changed (version
        counters); while a CUDA graph is being captured it ALWAYS runs, so that it becomes a node of the graph and every replay
        reads the masters the optimiser has just updated (a version check at capture time would freeze the capture-time images
        into every replay)."""
        if self._prec() != _cabi.PREC_BF16:
            return None
        dev = params[0].device
        if self._packed is None or self._packed.device != dev:
            self._packed = _ops.alloc_packed(dev)
            self._packed_version = None
        if self._pack_event is not None:          # begin_pack() already issued this step's pack on a side stream
            torch.cuda.current_stream().wait_event(self._pack_event)
            self._pack_event = None
            return self._packed
        capturing = torch.cuda.is_current_stream_capturing()
        version = tuple(p._version for p in params) + tuple(p.data_ptr() for p in params)
        if capturing or self._packed_version != version:
            _ops.pack_weights(params, out=self._packed)
            self._packed_version = None if capturing else version
        return self._packed

    def join_pack(self):
        """Join a pack that no render consumed (a step whose rgb / depth weights are both zero)."""
        if self._pack_event is not None:
            torch.cuda.current_stream().wait_event(self._pack_event)
            self._pack_event = None

    def begin_pack(self, stream):
        """Issue this step's weight pack on `stream` (forked from the current stream), off the critical path of the pose /
        ray-generation chain; the next `_packed_weights` call makes the render stream wait for it."""
        if self._prec() != _cabi.PREC_BF16:
            return
        params = self.model.param_list()
        if self._packed is None or self._packed.device != params[0].device:
            self._packed = _ops.alloc_packed(params[0].device)
        cur = torch.cuda.current_stream()
        stream.wait_stream(cur)
        with torch.cuda.stream(stream):
            _ops.pack_weights(params, out=self._packed)
            ev = torch.cuda.Event()
            ev.record(stream)
        self._packed_version = None
        self._pack_event = ev

    def nope_nerf(self, pixels, depth, camera_mat, world_mat, scale_mat, add_noise=False, it=100000, eval_=False,
                  noise=None, lazy_mask=False, shard=None):
        """shard = (lo, hi): data parallelism.  `pixels` / `depth` / `noise` hold the GLOBAL ray batch; ray
        generation (33 B/ray) runs on all of it on every rank, so that the validity mask and the depth targets of the whole
        batch are known everywhere without a collective, and only rays [lo, hi) go through the fused render.  The returned
        per-ray tensors are the shard's; `mask_global` / `depth_gt_global` cover the whole batch."""
        cfg = self.cfg
        ndc = cfg['sample_option'] == 'ndc'
        if cfg['sample_option'] not in ('uniform', 'ndc'):
            raise ValueError("unknown sample_option %r" % (cfg['sample_option'],))
        batch_size, n_points, _ = pixels.shape
        if batch_size != 1:
            raise ValueError("batch size 1 only (as the reference: model/rendering.py:80 takes element 0)")
        S = cfg['num_points'] - cfg['outside_steps']
        origin, dirs, nrm, d_gt, mask = _ops.RayGen.apply(pixels, depth, camera_mat, world_mat, scale_mat,
                                                          cfg['normalise_ray'])
        d_gt_global, mask_global = d_gt, mask
        if shard is not None:
            lo, hi = shard
            dirs, nrm, d_gt, mask = dirs[lo:hi], nrm[lo:hi], d_gt[lo:hi], mask[lo:hi]
            if noise is not None:
                noise = noise.reshape(n_points, S)[lo:hi]
            n_points = hi - lo
        ray_o = view = None
        dr = None
        dirs_world = dirs
        if ndc:
            # model/rendering.py:92-93,169-181: rays warped to NDC (near plane 1.0), samples on [0, 1] without jitter;
            # the view-direction encoding keeps the world-space ray.
            ray_o, ndc_d = _ops.NdcRays.apply(camera_mat, origin, dirs, 1.0)
            view, dirs = dirs, ndc_d
            add_noise, noise, dr = False, None, (0., 1.)
        if add_noise and noise is None:
            # same draw, shape and generator as the reference (model/rendering.py:190)
            noise = torch.rand(batch_size, n_points, S, device=pixels.device)
        if not add_noise:
            noise = None
        params = self.model.param_list()
        packed = self._packed_weights(params)
        need_grad = torch.is_grad_enabled() and (any(p.requires_grad for p in params) or dirs.requires_grad
                                                 or origin.requires_grad)
        if need_grad:
            outs = []
            step = MAX_TRAIN_RAYS_FP32 if self._prec() == _cabi.PREC_FP32 else n_points
            for r0 in range(0, n_points, step):
                r1 = min(n_points, r0 + step)
                rc = self._cfg(r1 - r0, S, dr)
                nz = None if noise is None else noise.reshape(n_points, S)[r0:r1]
                if ndc:
                    outs.append(_ops.RenderRaysEx.apply(rc, packed, ray_o[r0:r1], dirs[r0:r1], view[r0:r1], None, *params))
                else:
                    outs.append(_ops.RenderRays.apply(rc, packed, origin, dirs[r0:r1], nz, *params))
            rgb, dist, w, alpha, z = [torch.cat(t, 0) if len(outs) > 1 else t[0] for t in zip(*outs)]
        else:
            rc = self._cfg(n_points, S, dr)
            rgb, dist, w, alpha, z = _ops.render_nograd(rc, [p.detach() for p in params], packed, origin.detach(),
                                                        dirs.detach(), None if noise is None else noise.reshape(n_points, S),
                                                        ray_o=None if ray_o is None else ray_o.detach(),
                                                        view=None if view is None else view.detach())
        diff_norm = None
        if cfg['normal_loss'] and not eval_:
            # model/rendering.py:127-135: normals at the depth-map surface points and at jittered neighbours.  Nothing downstream
            # reads it (model/losses.py and model/training.py never touch out_dict['normal']), so it is computed forward only;
            # the draw comes after the stratified jitter, as in the reference.
            with torch.no_grad():
                d_i = d_gt.detach()
                dists = torch.where(torch.isfinite(d_i), d_i, torch.ones_like(d_i))
                dists = torch.where(d_i == 0, torch.zeros_like(dists), dists)
                points = origin.detach().view(1, 3) + dirs_world.detach() * dists.unsqueeze(-1)
                surface = points[mask]
                n_s = surface.shape[0]
                neigh = surface + (torch.rand_like(surface) - 0.5) * 0.01
                g = self.model.gradient(torch.cat([surface, neigh], dim=0), it)[:, 0, :]
                normals = g / (g.norm(2, dim=1).unsqueeze(-1) + 10 ** (-5))
                diff_norm = torch.norm(normals[:n_s] - normals[n_s:], dim=-1)
        if eval_ and cfg['normalise_ray']:  # model/rendering.py:144-148
            dist = dist / nrm
            d_gt = d_gt / nrm
        if not eval_ and not lazy_mask:
            # the reference's boolean indexing (model/rendering.py:151-153): data-dependent size => host sync.
            # Callers that want to stay asynchronous pass lazy_mask=True and use the dense tensors + `mask`.
            depth_pred, depth_gt = dist[mask], d_gt[mask]
        else:
            depth_pred, depth_gt = dist, d_gt
        if ndc:  # model/rendering.py:158-159
            depth_gt = 1 - 1 / depth_gt
            d_gt = 1 - 1 / d_gt
        return {
            'rgb': rgb.reshape(batch_size, -1, 3), 'z_vals': z, 'normal': diff_norm,
            'depth_pred': depth_pred, 'depth_gt': depth_gt, 'alpha': alpha,
            'depth_pred_dense': dist, 'depth_gt_dense': d_gt, 'mask': mask, 'weights': w,
            'depth_gt_global': d_gt_global, 'mask_global': mask_global,
        }
