"""Loss / Loss_Eval -- drop-in for the reference's model/losses.py.

forward() keeps the reference's signature, zero-weight skipping and its 10 return keys.  Every active term is a
kernel launch (rgb, depth l1 / invariant, Chamfer, reprojection photometric); the terms whose weights are [0,0] / False in
every shipped config (pose smoothness, depth consistency, t-cycle, the with_ssim mix of the photometric term) stay plain
PyTorch and are compared with the reference's own methods in tests/test_cpu_dropin.py.
Differences kept on purpose: tensors follow the inputs' device (no hard-coded .cuda()), and a NaN loss raises
FloatingPointError (Loss.raise_if_nan, called by the Trainer after every step, one step late so that the read never
stalls the launch queue) instead of dropping into a debugger (model/losses.py:213-214).
"""
import torch
from torch import nn
from torch.nn import functional as F

from . import _ops


class Loss_Eval(nn.Module):
    def forward(self, rgb_pred, rgb_gt):
        return {'loss': _ops.RgbLoss.apply(rgb_pred, rgb_gt)[2]}


class Loss(nn.Module):
    def __init__(self, cfg=None):
        super().__init__()
        self.depth_loss_type = cfg['depth_loss_type']
        self.cfg = cfg
        self.check_nan = True
        # data-parallel context (set by Trainer): the Chamfer SOURCE points are sharded across ranks
        self.dp = None
        self._nan_pending, self._nan_step = None, 0
        # names of the active terms -> [device vector, values last written].  ONE persistent device tensor
        # per active set: a captured graph bakes its address, the VALUES (the annealed weights change every epoch) are
        # written before each replay, so annealing does not multiply graphs
        self._weight_vectors = {}

    def weight_vector(self, names, values, dev):
        """Device vector of the active terms' weights.  Outside a capture the current values are (re)written when they changed
        (one small H2D copy); during a capture the tensor is used as it is -- Trainer.train_step_graphed
        writes the step's values through this same call before every replay."""
        rec = self._weight_vectors.get((names, str(dev)))
        vals = tuple(float(v) for v in values)
        if rec is None:
            rec = self._weight_vectors[(names, str(dev))] = [torch.tensor(vals, dtype=torch.float32, device=dev), vals]
        elif rec[1] != vals and not (torch.device(dev).type == 'cuda' and torch.cuda.is_current_stream_capturing()):
            rec[0].copy_(torch.tensor(vals, dtype=torch.float32))     # a few bytes, staged by the driver: stream-ordered before the next launch
            rec[1] = vals
        return rec[0]

    def raise_if_nan(self, loss_dict, no_valid_depth=None):
        """model/losses.py:213-214 (`if torch.isnan(loss): breakpoint()`) without a host synchronisation on the step just
        issued: the loss is copied to pinned host memory asynchronously and examined when the NEXT step has been queued.
        `no_valid_depth` (optional device scalar, Trainer's graphed step with sparse depth): non-zero when neither of the step's
        two ray draws contained a pixel with a valid depth -- the case the reference's loop resamples until it is gone
        (model/training.py:279-284); reported the same way, one step late."""
        loss = loss_dict['loss']
        if not loss.is_cuda:
            if bool(torch.isnan(loss)):
                raise FloatingPointError("NaN loss")
            return
        prev = self._nan_pending
        host = prev[1] if prev is not None else torch.zeros(2, 2, dtype=torch.float32).pin_memory()
        slot = 0
        if prev is not None:
            ev, _, slot, step = prev
            ev.synchronize()
            if bool(torch.isnan(host[slot, 0])):
                self._nan_pending = None
                raise FloatingPointError("NaN loss at step %d (model/losses.py:213)" % step)
            if float(host[slot, 1]) != 0.0:
                self._nan_pending = None
                raise RuntimeError("step %d: no ray with a valid depth in two consecutive draws (model/training.py:279-284 resamples "
                                   "until there is one); run this batch kernel by kernel (Trainer.use_graph = False)" % step)
            slot ^= 1
        host[slot, 0:1].copy_(loss.detach().reshape(1), non_blocking=True)
        if no_valid_depth is not None:
            host[slot, 1:2].copy_(no_valid_depth.detach().reshape(1).float(), non_blocking=True)
        else:
            host[slot, 1] = 0.0
        ev = torch.cuda.Event()
        ev.record()
        self._nan_pending = (ev, host, slot, self._nan_step)
        self._nan_step += 1

    # -- individual terms (reference method names) ----------------------------------------------
    def get_rgb_full_loss(self, rgb_values, rgb_gt, rgb_loss_type='l2'):
        out = _ops.RgbLoss.apply(rgb_values, rgb_gt)
        return out[0] if rgb_loss_type == 'l1' else out[1]

    def depth_loss_dpt(self, pred_depth, gt_depth, weight=None, mask=None):
        if weight is not None:
            raise NotImplementedError("weighted invariant depth loss is never called by the reference")
        return _ops.DepthInvariant.apply(pred_depth, gt_depth, mask)

    def get_depth_loss(self, depth_pred, depth_gt, mask=None):
        if self.depth_loss_type == 'l1':
            return _ops.DepthL1.apply(depth_pred, depth_gt, mask)
        if self.depth_loss_type == 'invariant':
            return self.depth_loss_dpt(depth_pred, depth_gt, mask=mask)
        raise ValueError("unknown depth_loss_type %r" % (self.depth_loss_type,))

    def comp_closest_pts_idx_with_split(self, pts_src, pts_des):
        """(3,S),(3,D) -> (S,) int64, like the reference; no (3,S,D) tensor is ever built."""
        return _ops.chamfer_nn(pts_src.t(), pts_des.t())

    def comp_point_point_error(self, Xt, Yt):
        """Xt (3,S), Yt (3,D)."""
        X, Y = Xt.t().contiguous(), Yt.t().contiguous()
        return _ops.ChamferErr.apply(X, Y, _ops.chamfer_nn(X, Y))

    def _point_point_error_rows(self, X, Y):
        """mean_i |X_i - Y[nn(i)]| over the rows of X (P,3).  With world_size G > 1 this rank searches only its slice
        of the source rows (both clouds stay whole on every rank) and returns G * (its partial sum) / P: the gradient
        allreduce AVERAGES over ranks, so the average of these terms -- value and gradient -- is the full-batch one."""
        dp = self.dp
        if dp is None or dp.world_size == 1:
            return _ops.ChamferErr.apply(X, Y, _ops.chamfer_nn(X, Y))
        n = X.shape[0]
        lo, hi = dp.shard(n)
        if hi == lo:
            return X.sum() * 0.0
        Xs = X[lo:hi]
        return _ops.ChamferErr.apply(Xs, Y, _ops.chamfer_nn(Xs, Y)) * (float(hi - lo) * dp.world_size / float(n))

    def get_pc_loss(self, Xt, Yt):
        if self.cfg['match_method'] != 'dense':
            raise ValueError("match_method %r" % (self.cfg['match_method'],))
        X, Y = Xt[0], Yt[0]  # (P,3) row-major: the layout the kernels want (the reference permutes to (3,P))
        return self._point_point_error_rows(X, Y) + self._point_point_error_rows(Y, X)

    def mean_on_mask(self, diff, valid_mask):
        """model/losses.py:79-87: mean of `diff` over the masked elements, 0 for an empty mask -- without the host read."""
        mask = valid_mask.expand_as(diff)
        return (diff * mask).sum() / mask.sum().clamp_min(1)

    def get_rgb_s_loss(self, rgb1, rgb2, valid_points):
        if self.cfg['with_ssim']:
            # False in every shipped config: plain PyTorch, model/losses.py:152-159 with the SSIM branch
            diff = 0.15 * (rgb1 - rgb2).abs().clamp(0, 1) + 0.85 * compute_ssim_loss(rgb1, rgb2)
            return self.mean_on_mask(diff, valid_points)
        return _ops.Photometric.apply(rgb1, rgb2, valid_points)

    # zero-weight terms: plain PyTorch, same formulas as model/losses.py:105-128,161-162
    def get_weight_dist_loss(self, t_list):
        dist = (t_list - t_list.roll(shifts=1, dims=0))[1:].norm(dim=1)
        dist_diff = (dist - dist.roll(shifts=1))[1:]
        return dist.mean(), dist_diff.pow(2.0).mean()

    def get_depth_consistency_loss(self, d1_proj, d2, d2_proj=None, d1=None):
        loss = (d1_proj - d2).abs().sum() / float(d1_proj.shape[1])
        if d2_proj is not None:
            loss = 0.5 * loss + 0.5 * (d2_proj - d1).abs().sum() / float(d2_proj.shape[1])
        return loss

    def get_t_cycle_loss(self, rt_pred, rt_gt):
        return torch.norm(torch.eye(4, device=rt_gt.device) - torch.inverse(rt_gt) @ rt_pred)

    # -- aggregate --------------------------------------------------------------------------------
    def forward(self, rgb_pred, rgb_gt, depth_pred=None, depth_gt=None, t_list=None, X=None, Y=None, rgb_pc1=None,
                rgb_pc1_proj=None, valid_points=None, d1_proj=None, d2=None, d2_proj=None, d1=None, weights={},
                rgb_loss_type='l2', depth_mask=None, **kwargs):
        dev = rgb_gt.device
        zero = torch.zeros((), device=dev)
        # Data parallelism: the rendered tensors are this rank's shard [lo, hi) of the global batch and the
        # gradient allreduce AVERAGES over the G ranks, so every per-ray term is returned as G x (its share of the GLOBAL
        # normaliser): the rank average of the values is the 1-GPU loss and the averaged gradient is the 1-GPU gradient,
        # also for unequal shards and ragged masks.
        dpi = kwargs.get('dp_info')
        G = self.dp.world_size if (dpi is not None and self.dp is not None) else 1
        rgb3 = None
        if weights['rgb_weight'] != 0.0 or weights['depth_weight'] != 0.0:
            rgb3 = _ops.RgbLoss.apply(rgb_pred, rgb_gt)            # sums over the shard / shard size
            if dpi is not None:
                rgb3 = rgb3 * (float(G) * float(dpi['hi'] - dpi['lo']) / float(dpi['n_global']))
        rgb_full_loss = (rgb3[0] if rgb_loss_type == 'l1' else rgb3[1]) if weights['rgb_weight'] != 0.0 else zero
        if weights['depth_weight'] == 0.0:
            depth_loss = zero
        elif dpi is None:
            depth_loss = self.get_depth_loss(depth_pred, depth_gt, depth_mask)
        elif self.depth_loss_type == 'l1':
            # model/losses.py:60-63 divides by the number of valid rays of the WHOLE batch (known on every rank: mask_global)
            c_loc = depth_mask.sum().float() if depth_mask is not None else torch.tensor(float(depth_pred.numel()), device=dev)
            c_glob = dpi['mask_global'].sum().float().clamp_min(1.0)
            depth_loss = self.get_depth_loss(depth_pred, depth_gt, depth_mask) * (c_loc * float(G) / c_glob)
        else:
            # model/losses.py:35-58 takes the median and the mean absolute deviation of the WHOLE batch: all-gather the rendered
            # depths (R x 4 B), evaluate the term redundantly on every rank, keep the gradient of this rank's rays only
            from . import parallel
            lo, hi = dpi['lo'], dpi['hi']
            pred_all = parallel.AllGatherRows.apply(depth_pred, self.dp, dpi['n_global'])
            g_all = dpi['depth_gt_global']
            gt_all = torch.cat([g_all[:lo].detach(), depth_gt, g_all[hi:].detach()])
            full = self.get_depth_loss(pred_all, gt_all, dpi['mask_global'])
            # every rank evaluates the SAME global term; a rank keeps only its own rays' slice of the gradient (AllGatherRows),
            # so the averaged gradient needs G x that slice, while the VALUE must stay L (the rank mean of L is L):
            depth_loss = float(G) * full - float(G - 1) * full.detach()
        if weights['weight_dist_2nd_loss'] != 0.0 or weights['weight_dist_1st_loss'] != 0.0:
            loss_dist_1st, loss_dist_2nd = self.get_weight_dist_loss(t_list)
        else:
            loss_dist_1st, loss_dist_2nd = zero, zero
        # 'pc_loss_pre' / 'rgb_s_loss_pre': the same two calls, already made by the trainer on its reference-frame stream;
        # 'join_pre' makes the current stream wait for that stream -- only here, so that the terms above do not wait for it
        if kwargs.get('join_pre') is not None:
            kwargs['join_pre']()
        pc_loss, rgb_s_loss = kwargs.get('pc_loss_pre'), kwargs.get('rgb_s_loss_pre')
        if pc_loss is None:
            pc_loss = self.get_pc_loss(X, Y) if weights['pc_weight'] != 0.0 else zero
        if rgb_s_loss is None:
            rgb_s_loss = self.get_rgb_s_loss(rgb_pc1, rgb_pc1_proj, valid_points) if weights['rgb_s_weight'] != 0.0 else zero
        if weights['depth_consistency_weight'] != 0.0:
            depth_consistency_loss = self.get_depth_consistency_loss(d1_proj, d2, d2_proj, d1)
        else:
            depth_consistency_loss = zero
        t_cycle_loss = self.get_t_cycle_loss(kwargs['rt_12'], kwargs['rt_12_gt']) if weights['t_cycle_weight'] != 0.0 else zero
        rgb_l2_mean = rgb3[2] if rgb3 is not None else zero

        # sum_k weight_k * term_k over the active terms (model/losses.py:204-208) as one stack + one dot instead of a chain of
        # scalar multiplies and adds: two launches forward, one backward
        named = (('rgb_weight', rgb_full_loss), ('depth_weight', depth_loss), ('weight_dist_1st_loss', loss_dist_1st),
                 ('weight_dist_2nd_loss', loss_dist_2nd), ('pc_weight', pc_loss), ('rgb_s_weight', rgb_s_loss),
                 ('depth_consistency_weight', depth_consistency_loss), ('t_cycle_weight', t_cycle_loss))
        active = [(k, float(weights[k]), v) for k, v in named if weights[k] != 0.0]
        if active:
            wvec = self.weight_vector(tuple(k for k, _, _ in active), [w for _, w, _ in active], dev)
            loss = torch.dot(torch.stack([v.reshape(()) for _, _, v in active]), wvec)
        else:
            loss = zero
        return {
            'loss': loss, 'loss_rgb': rgb_full_loss, 'loss_depth': depth_loss, 'l2_mean': rgb_l2_mean,
            'loss_dist_1st': loss_dist_1st, 'loss_dist_2nd': loss_dist_2nd, 'loss_pc': pc_loss,
            'loss_rgb_s': rgb_s_loss, 'loss_depth_consistency': depth_consistency_loss, 'loss_t_cycle': t_cycle_loss,
        }


class SSIM(nn.Module):
    """model/losses.py:232-261: (1 - SSIM) / 2 on 3 x 3 box windows after a one-pixel reflection pad, clamped to [0, 1].  The
    reference feeds it the (B, h, w, 3) reprojection tensors as they are, i.e. pooled over its last two axes; so does this."""
    C1, C2 = 0.01 ** 2, 0.03 ** 2

    def forward(self, x, y):
        x, y = F.pad(x, (1, 1, 1, 1), mode='reflect'), F.pad(y, (1, 1, 1, 1), mode='reflect')
        box = lambda t: F.avg_pool2d(t, 3, 1)
        mx, my = box(x), box(y)
        vx, vy, cxy = box(x * x) - mx * mx, box(y * y) - my * my, box(x * y) - mx * my
        num = (2 * mx * my + self.C1) * (2 * cxy + self.C2)
        den = (mx * mx + my * my + self.C1) * (vx + vy + self.C2)
        return ((1 - num / den) / 2).clamp(0, 1)


compute_ssim_loss = SSIM()
