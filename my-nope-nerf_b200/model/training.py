"""Trainer -- drop-in for the reference's model/training.py (same constructor, train_step, compute_loss,
process_data_dict, anneal; same loss_dict keys plus `scale`, `shift`).

compute_loss keeps the reference's control flow (weight annealing, reference-frame pairing, zero-weight skipping,
RNG draw order: the ray choice, then rand for the stratified jitter) and replaces its tensor math with the
C-ABI kernels:  pose exp-map, 4x4 inverse, distortion, fused ray gather, ray generation, fused render,
point-cloud back-projection, reprojection sampling and the loss kernels.  Multi-GPU: every rank draws the same
global ray batch and takes its contiguous shard; gradients are averaged with ONE flat NCCL allreduce (parallel.py).
"""
import contextlib
import logging
import os

import numpy as np
import torch

from . import _cabi, _ops, parallel
from .extracting_images import _imwrite
from .losses import Loss

logger_py = logging.getLogger(__name__)

WEIGHT_NAMES = ['rgb_weight', 'depth_weight', 'pc_weight', 'rgb_s_weight', 'depth_consistency_weight',
                'weight_dist_2nd_loss', 'weight_dist_1st_loss', 't_cycle_weight']
# the order in which Loss.forward sums the terms (model/losses.py:204-208): the key of Loss.weight_vector
WEIGHT_NAMES_IN_SUM = ['rgb_weight', 'depth_weight', 'weight_dist_1st_loss', 'weight_dist_2nd_loss', 'pc_weight', 'rgb_s_weight',
                       'depth_consistency_weight', 't_cycle_weight']


def _prefer_fused_adam(opt):
    """train.py:59,100,118,141 build plain `torch.optim.Adam(params, lr=...)`: torch then steps with its multi-tensor `foreach` path,
    seven launches per optimiser (160 us for the 24 MLP tensors: a tenth of a 1,024-ray step).  The same class has a one-launch
    `fused` implementation of the same update; it is selected here, on the caller's optimiser, when nothing the caller chose
    stands against it (CUDA fp32 parameters, `foreach` / `fused` left at their defaults, none of amsgrad / maximize / capturable /
    differentiable).  Step counters already restored from a checkpoint move to the parameters' device, as the fused path needs.
    NNB_FUSED_TORCH_ADAM=0 leaves the optimisers untouched."""
    if type(opt) is not torch.optim.Adam or os.environ.get('NNB_FUSED_TORCH_ADAM', '1') == '0':
        return
    for g in opt.param_groups:
        if g.get('fused', 0) is not None or g.get('foreach') is not None:
            return                                   # option missing in this torch, or set by the caller
        if g.get('amsgrad') or g.get('maximize') or g.get('capturable') or g.get('differentiable'):
            return
        if not all(p.is_cuda and p.dtype == torch.float32 for p in g['params']):
            return
    for g in opt.param_groups:
        g['fused'] = True
        for p in g['params']:
            st = opt.state.get(p)
            if st and torch.is_tensor(st.get('step')):
                st['step'] = st['step'].to(device=p.device, dtype=torch.float32)


class Trainer(object):
    def __init__(self, model, optimizer, cfg, device=None, optimizer_pose=None, pose_param_net=None,
                 optimizer_focal=None, focal_net=None, optimizer_distortion=None, distortion_net=None, **kwargs):
        self.model = model
        self.optimizer = optimizer
        self.device = device
        for opt in (optimizer, optimizer_pose, optimizer_focal, optimizer_distortion):
            if opt is not None:
                _prefer_fused_adam(opt)
        self.fork_wgrad = True      # train_step_graphed: weight-gradient kernel on its own stream (joined before the optimiser)
        self.fork_geometry = True   # train_step_graphed: reference-frame branch as a parallel branch of the graph
        self._geom_stream = None    # set while train_step_graphed warms up and captures: forks the reference-frame branch
        self._geom_forked = False
        self._graph_body = False    # True while train_step_graphed warms up and captures (graph-safe forms of host-side decisions)
        self._no_valid_depth = None
        self.optimizer_pose = optimizer_pose
        self.pose_param_net = pose_param_net
        self.focal_net = focal_net
        self.optimizer_focal = optimizer_focal
        self.distortion_net = distortion_net
        self.optimizer_distortion = optimizer_distortion

        self.n_training_points = cfg['n_training_points']
        self.rendering_technique = cfg['type']
        self.vis_geo = cfg['vis_geo']
        self.detach_gt_depth = cfg['detach_gt_depth']
        self.pc_ratio = cfg['pc_ratio']
        self.match_method = cfg['match_method']
        self.shift_first = cfg['shift_first']
        self.detach_ref_img = cfg['detach_ref_img']
        self.scale_pcs = cfg['scale_pcs']
        self.detach_rgbs_scale = cfg['detach_rgbs_scale']
        self.vis_reprojection_every = cfg['vis_reprojection_every']
        self.nearest_limit = cfg['nearest_limit']
        self.annealing_epochs = cfg['annealing_epochs']
        for w in WEIGHT_NAMES:
            setattr(self, w, cfg[w])
        self.loss = Loss(cfg)
        # data-parallel context (world size 1 unless torch.distributed is initialised)
        self.dp = kwargs.get('dp') or parallel.DataParallelContext.from_env()
        self.loss.dp = self.dp
        # test hooks: inject the ray choice / jitter instead of drawing them (parity tests)
        self.fixed_ray_idx = None
        self.fixed_noise = None
        # 'auto' (default; env NNB_GRAPH=0/1 overrides): `train_step`, the call train.py:213 makes, replays the step as a CUDA
        # graph whenever the batch allows it (see _graph_ok); False: always launch kernel by kernel; True: always graph
        env = os.environ.get('NNB_GRAPH', 'auto')
        self.use_graph = {'0': False, '1': True}.get(env, 'auto')
        self.max_graphs = int(os.environ.get('NNB_MAX_GRAPHS', '8'))   # captured step variants kept (least recently used evicted)
        # 'feistel' (nnb_ray_choice, no sort) or 'reference' (torch.randperm(h*w)[:n], the draw of model/training.py:277 with
        # the same consumption of the device generator); cfg key `ray_choice`, env NNB_RAY_CHOICE
        self.ray_choice = cfg.get('ray_choice', os.environ.get('NNB_RAY_CHOICE', 'feistel'))
        if self.ray_choice not in ('feistel', 'reference'):
            raise ValueError("ray_choice must be 'feistel' or 'reference' (got %r)" % (self.ray_choice,))

    # -- reference API -------------------------------------------------------------------------
    def train_step(self, data, it=None, epoch=None, scheduling_start=None, render_path=None):
        """model/training.py:70-100.  Same signature and return value; on a CUDA device the step is replayed as a CUDA graph
        (train_step_graphed) when the batch qualifies, else launched kernel by kernel (train_step_eager).  Both run the same
        kernels; tests/test_gpu_step.py checks that they produce the same losses and parameters."""
        if self._graph_ok(data):
            return self.train_step_graphed(data, it=it, epoch=epoch, scheduling_start=scheduling_start)
        return self.train_step_eager(data, it=it, epoch=epoch, scheduling_start=scheduling_start, render_path=render_path)

    def _graph_ok(self, data):
        if self.use_graph is False:
            return False
        if self.use_graph is True:
            return True
        dev = torch.device(self.device) if self.device is not None else None
        return (dev is not None and dev.type == 'cuda' and self.pose_param_net is not None
                and self.fixed_ray_idx is None and self.fixed_noise is None
                and self.ray_choice != 'reference'        # the reference's generator sequence is an eager-mode property
                and not self.optimizer_focal              # learned intrinsics (non-default): kernel-by-kernel step
                and _cabi.profile is None)                # per-entry-point timing (bench.py) needs eager launches

    def train_step_eager(self, data, it=None, epoch=None, scheduling_start=None, render_path=None):
        self.model.train()
        self.optimizer.zero_grad()
        if self.pose_param_net:
            self.pose_param_net.train()
            self.optimizer_pose.zero_grad()
        if self.focal_net:
            self.focal_net.train()
            self.optimizer_focal.zero_grad()
        if self.distortion_net:
            self.distortion_net.train()
            self.optimizer_distortion.zero_grad()
        loss_dict = self.compute_loss(data, it=it, epoch=epoch, scheduling_start=scheduling_start,
                                      out_render_path=render_path)
        loss_dict['loss'].backward()
        if self.dp.world_size > 1:
            nets = [self.model, self.pose_param_net, self.distortion_net, self.focal_net]
            self.dp.allreduce_grads([p for n in nets if n is not None for p in n.parameters()], extra=self._loss_scalars(loss_dict))
        self.optimizer.step()
        if self.optimizer_pose:
            self.optimizer_pose.step()
        if self.optimizer_focal:
            self.optimizer_focal.step()
        if self.optimizer_distortion:
            self.optimizer_distortion.step()
        if self.loss.check_nan:
            self.loss.raise_if_nan(loss_dict)
        return loss_dict

    # -- the same step as a CUDA graph -----------------------------------------
    def train_step_graphed(self, data, it=None, epoch=None, scheduling_start=None):
        """`train_step` with compute_loss + backward captured in a CUDA graph and replayed: one launch instead of ~150, no
        launch gaps on the device.  Everything that varies from step to step flows through device memory -- the batch is
        copied into static input buffers, the frame indices travel as device tensors (LearnPose / Learn_Distortion gather
        their rows on the device), ray choice and jitter are drawn inside the graph (graph-safe Philox state).  What stays
        host-side selects WHICH graph runs: the last-frame branch of model/training.py:322-345, the fix_scaleN pin, the
        annealed loss weights and the tensor shapes; a new combination is captured on first use (two eager warm-up passes
        + capture, ~50 ms).  The optimiser steps and the gradient allreduce run eagerly after the replay.
        Returns the loss dict of the replayed graph (static tensors, overwritten by the next replay of the same graph)."""
        if not hasattr(self, '_graphs'):
            self._graphs = {}
        n = self.pose_param_net.num_cams
        idx = int(data.get('img.idx'))
        ref = int(data.get('img.ref_idxs')) if data.get('img.ref_idxs') is not None else -1
        wvals = [self.anneal(getattr(self, w)[0], getattr(self, w)[1], scheduling_start, self.annealing_epochs, epoch) for w in WEIGHT_NAMES]
        # WHICH terms are active selects the graph (zero-weight terms are skipped, model/losses.py:169-197); their VALUES -- the
        # annealed weights change every epoch -- live in a device vector written below, before the replay
        wkey = tuple(v != 0.0 for v in wvals) + (epoch < self.annealing_epochs + scheduling_start,)
        # the tensors the step reads (the DataLoader also yields img.gt_depths, evaluation only: not staged)
        tensors = sorted(k for k, v in data.items() if torch.is_tensor(v) and k in self.STEP_KEYS and k not in ('img.idx', 'img.ref_idxs'))
        # injected draws (test hooks) are baked into a capture as device pointers: they select their own graphs
        hooks = tuple(None if t is None else (t.data_ptr(), str(t.device)) for t in (self.fixed_ray_idx, self.fixed_noise))
        key = (idx < n - 1, idx == n - 1, ref == n - 1, wkey, hooks, tuple((k, tuple(data[k].shape), data[k].dtype) for k in tensors))
        slot = self._graphs.pop(key, None)
        if slot is None:
            while len(self._graphs) >= self.max_graphs:       # each graph owns a saved-activation workspace (1.3 MB per ray)
                self._graphs.pop(next(iter(self._graphs)))     # least recently used first (dicts keep insertion order)
            slot = self._capture_step(data, tensors, it, epoch, scheduling_start)
        self._graphs[key] = slot                              # most recently used last
        wnow = dict(zip(WEIGHT_NAMES, wvals))
        active = tuple(w for w in WEIGHT_NAMES_IN_SUM if wnow[w] != 0.0)
        if active:
            self.loss.weight_vector(active, [wnow[w] for w in active], self.device)    # this step's weight values, before the replay
        static = slot['static']
        # batch staging: every device-resident tensor of the batch and the two frame indices in ONE launch (nnb_copy_many)
        pairs = []
        for k in tensors:
            src, dst = data[k], static[k]
            if src.is_cuda and src.is_contiguous() and src.dtype == dst.dtype and src.device == dst.device:
                pairs.append((dst, src))
            else:
                dst.copy_(src, non_blocking=True)
        scalars = [(static['img.idx_dev'], idx)]
        if ref >= 0:
            scalars.append((static['img.ref_idxs_dev'], ref))
        _ops.copy_many(pairs, scalars)
        static['img.idx'], static['img.ref_idxs'] = data.get('img.idx'), data.get('img.ref_idxs')
        slot['graph'].replay()
        self._last_draw = slot['last_draw']      # this variant's ray choice / jitter tensors (parity tests read them)
        self.graphed_launches = getattr(self, 'graphed_launches', 0) + slot['n_launches']
        for p in slot['no_grad']:
            p.grad = None       # a parameter this graph variant gives no gradient must not carry one from another variant
        for p, g in slot['grads']:
            p.grad = g
        self._optimizer_tail(slot['loss_dict'])
        if self.loss.check_nan:
            self.loss.raise_if_nan(slot['loss_dict'], slot['no_valid_depth'])
        return slot['loss_dict']

    # keys of the batch schema the training step reads (process_data_dict / process_data_reference); the DataLoader also yields
    # `img.gt_depths` (evaluation only, model/training.py never moves it to the device)
    STEP_KEYS = ('img', 'img.idx', 'img.dpt', 'img.depth', 'img.depth_mask', 'img.camera_mat', 'img.scale_mat', 'img.pose_gt',
                 'img.ref_imgs', 'img.ref_dpts', 'img.ref_depths', 'img.ref_idxs', 'img.ref_pose_gt')

    def upload_batch(self, data, stream=None):
        """Host batch (pinned, as the DataLoader yields it) -> dict with the tensors the step reads on the device, copied
        asynchronously on `stream` (default: the current stream); the two frame indices stay on the host.  Returns
        (device_batch, event) -- make the compute stream wait on the event before the step."""
        stream = stream or torch.cuda.current_stream()
        with torch.cuda.stream(stream):
            out = {k: (v.to(self.device, non_blocking=True) if torch.is_tensor(v) and k not in ('img.idx', 'img.ref_idxs') else v)
                   for k, v in data.items() if k in self.STEP_KEYS}
            ev = torch.cuda.Event()
            ev.record(stream)
        return out, ev

    def _all_nets(self):
        return [m for m in (self.model, self.pose_param_net, self.distortion_net, self.focal_net) if m is not None]

    LOSS_KEYS = ('loss', 'loss_rgb', 'loss_depth', 'l2_mean', 'loss_dist_1st', 'loss_dist_2nd', 'loss_pc', 'loss_rgb_s',
                 'loss_depth_consistency', 'loss_t_cycle')

    def _loss_scalars(self, loss_dict):
        """The loss terms the gradient allreduce averages along with the gradients: a rank's per-ray terms are G x its share
        of the global normaliser (Loss.forward), so their rank average is the 1-GPU value.  Distinct storages only."""
        seen, out = set(), []
        for k in self.LOSS_KEYS:
            v = loss_dict.get(k)
            if torch.is_tensor(v) and v.numel() == 1 and v.data_ptr() not in seen:
                seen.add(v.data_ptr())
                out.append(v)
        return out

    def _optimizer_tail(self, loss_dict=None):
        if self.dp.world_size > 1:
            self.dp.allreduce_grads([p for m in self._all_nets() for p in m.parameters()],
                                    extra=self._loss_scalars(loss_dict) if loss_dict is not None else ())
        for opt in (self.optimizer, self.optimizer_pose, self.optimizer_focal, self.optimizer_distortion):
            if opt:
                opt.step()

    def _capture_step(self, data, tensors, it, epoch, scheduling_start):
        dev = self.device
        for m in self._all_nets():
            m.train()
        static = {k: data[k].to(dev, copy=True) for k in tensors}
        static['img.idx_dev'] = data['img.idx'].reshape(1).to(dev, dtype=torch.int64)
        static['img.idx'] = data.get('img.idx')
        static['img.ref_idxs'] = data.get('img.ref_idxs')
        if data.get('img.ref_idxs') is not None:
            static['img.ref_idxs_dev'] = data['img.ref_idxs'].reshape(1).to(dev, dtype=torch.int64)
        params = [p for m in self._all_nets() for p in m.parameters() if p.requires_grad]

        geom = torch.cuda.Stream() if self.fork_geometry else None
        # the weight-gradient kernel of the render backward runs on its own stream: it leaves room on every SM (64 registers x
        # 192 threads), so the ray-generation / pose backward chain that follows the data-gradient kernel runs beside it
        wgrad = torch.cuda.Stream() if (self.fork_wgrad and self.model.renderer.precision == 'bf16') else None

        renderer = self.model.renderer
        pack = torch.cuda.Stream() if renderer.precision == 'bf16' else None

        def body():
            for p in params:
                p.grad = None
            self._geom_stream, self._geom_forked = geom, False
            self._graph_body, self._no_valid_depth = True, None
            _ops.wgrad_stream, _ops.wgrad_forked = wgrad, False
            try:
                if pack is not None:
                    # the bf16 weight images are re-packed from the fp32 masters INSIDE the graph (every replay follows an
                    # optimiser step), on a side stream beside the pose / ray-generation chain
                    renderer.begin_pack(pack)
                ld = self.compute_loss(static, it=it, epoch=epoch, scheduling_start=scheduling_start)
                ld['loss'].backward()
                renderer.join_pack()                                # no-op when the render consumed the pack
                if geom is not None and self._geom_forked:          # (a step without reference-frame terms never forks it)
                    torch.cuda.current_stream().wait_stream(geom)   # nothing of the branch is left unjoined when capture ends
                if wgrad is not None and _ops.wgrad_forked:
                    torch.cuda.current_stream().wait_stream(wgrad)  # parameter gradients complete before the optimiser reads them
            finally:
                self._geom_stream = None
                self._graph_body = False
                _ops.wgrad_stream = None
            return ld

        warn = getattr(torch.autograd.graph, 'set_warn_on_accumulate_grad_stream_mismatch', None)
        if warn is not None:
            warn(False)   # warm-up and capture run on side streams on purpose
        cur = torch.cuda.current_stream()
        side = torch.cuda.Stream()
        side.wait_stream(cur)
        with torch.cuda.stream(side):      # eager warm-up passes (allocator, lazily built tables); gradients are discarded
            body()
            body()
        cur.wait_stream(side)
        torch.cuda.synchronize()
        # the warm-up passes left their saved-activation workspace (1.3 MB per ray) cached in the ordinary pool; the capture
        # allocates its own in the graph's private pool -- hand the cached blocks back first (matters at 65,536 rays: 86 GB)
        torch.cuda.empty_cache()
        graph = torch.cuda.CUDAGraph()
        for p in params:
            p.grad = None
        l0 = _cabi.kernel_launches()
        with torch.cuda.graph(graph):
            loss_dict = body()
        return {'graph': graph, 'static': static, 'loss_dict': loss_dict, 'n_launches': _cabi.kernel_launches() - l0,
                'no_valid_depth': self._no_valid_depth, 'last_draw': getattr(self, '_last_draw', None),
                'grads': [(p, p.grad) for p in params if p.grad is not None],
                'no_grad': [p for p in params if p.grad is None]}

    def process_data_dict(self, data):
        device = self.device
        img = data.get('img').to(device)
        img_idx = data.get('img.idx')
        depth = data.get('img.dpt')
        if depth is None:
            depth = data.get('img.depth')
        depth = depth.to(device).unsqueeze(1)
        depth_mask = data.get('img.depth_mask').to(device).unsqueeze(1)
        camera_mat = data.get('img.camera_mat').to(device)
        scale_mat = data.get('img.scale_mat').to(device)
        pose_gt = data.get('img.pose_gt').to(device)
        return (img, depth, camera_mat, scale_mat, img_idx, pose_gt, depth_mask)

    def process_data_reference(self, data):
        device = self.device
        ref_imgs = data.get('img.ref_imgs').to(device)
        ref_depths = data.get('img.ref_dpts')
        if ref_depths is None:
            ref_depths = data.get('img.ref_depths')
        ref_depths = ref_depths.to(device).unsqueeze(1)
        ref_pose_gt = data.get('img.ref_pose_gt').to(device)
        ref_idxs = data.get('img.ref_idxs')
        return (ref_imgs, ref_depths, ref_idxs, ref_pose_gt)

    def anneal(self, start_weight, end_weight, anneal_start_epoch, anneal_epoches, current):
        if current <= anneal_start_epoch:
            return start_weight
        elif current >= anneal_start_epoch + anneal_epoches:
            return end_weight
        return start_weight + (end_weight - start_weight) * (current - anneal_start_epoch) / anneal_epoches

    # -- the step ---------------------------------------------------------------------------------
    def compute_loss(self, data, eval_mode=False, it=None, epoch=None, scheduling_start=None, out_render_path=None):
        weights = {w: self.anneal(getattr(self, w)[0], getattr(self, w)[1], scheduling_start, self.annealing_epochs, epoch)
                   for w in WEIGHT_NAMES}
        rgb_loss_type = 'l1' if epoch < self.annealing_epochs + scheduling_start else 'l2'
        render_model = (weights['rgb_weight'] != 0.0) or (weights['depth_weight'] != 0.0)
        use_ref_imgs = (weights['pc_weight'] != 0.0) or (weights['rgb_s_weight'] != 0.0) or (weights['t_cycle_weight'] != 0.0)

        n_points = self.n_training_points
        nl = self.nearest_limit
        (img, depth_raw, camera_mat_gt, scale_mat, img_idx, pose_gt, depth_mask) = self.process_data_dict(data)
        if use_ref_imgs:
            (ref_img, depth_ref_raw, ref_idx, ref_pose_gt) = self.process_data_reference(data)
        device = self.device
        batch_size, _, h, w = img.shape
        _, _, h_depth, w_depth = depth_raw.shape
        kwargs = dict()
        kwargs['t_list'] = self.pose_param_net.get_t()
        kwargs['weights'] = weights
        kwargs['rgb_loss_type'] = rgb_loss_type
        img_idx_i = int(img_idx)
        idx_dev, ref_idx_dev = data.get('img.idx_dev'), data.get('img.ref_idxs_dev')   # set by train_step_graphed only

        num_cams = self.pose_param_net.num_cams
        c2w = self.pose_param_net(img_idx_i, idx_dev)              # exp-map kernel
        world_mat = _ops.Mat4Inverse.apply(c2w).unsqueeze(0)       # model/training.py:257

        def affine(ss):
            """The kernels apply depth * a + b.  shift_first (model/training.py:261-264,312-315) means (depth + shift) * scale
            = depth * scale + shift * scale: same kernels with b = shift * scale (gradients through the two torch ops)."""
            if ss is None or not self.shift_first:
                return ss
            return torch.stack([ss[0], ss[0] * ss[1]])

        ss_raw = ss_in = None
        if self.distortion_net is not None:
            ss_raw = self.distortion_net.scale_shift(img_idx_i, idx_dev)   # [scale, shift] on device, no host sync
            ss_in = affine(ss_raw)
        if self.optimizer_focal:
            fxfy = self.focal_net(0)
            pad = torch.zeros(4, device=device)
            one = torch.ones(1, device=device)
            camera_mat = torch.cat([fxfy[0:1], pad, -fxfy[1:2], pad, -one, pad, one]).view(1, 4, 4)
        else:
            camera_mat = camera_mat_gt

        # The reference-frame branch (point clouds, Chamfer, reprojection: model/training.py:318-372) depends on the depth maps,
        # poses and distortion parameters only, not on the rendered rays.  It is issued BEFORE the ray choice and the render so that, when a second
        # stream is set (train_step_graphed), its many small launches become a parallel branch of the captured graph and run
        # beside the persistent render kernels, in the forward and (autograd replays each node on its forward stream) the
        # backward pass.  Neither branch draws random numbers, so the RNG order (ray choice, then rand) is unchanged.
        geom = self._geom_stream if use_ref_imgs else None
        if use_ref_imgs:
            main_stream = torch.cuda.current_stream() if geom is not None else None
            if geom is not None:
                geom.wait_stream(main_stream)
                self._geom_forked = True
            with (torch.cuda.stream(geom) if geom is not None else contextlib.nullcontext()):
                ref_idx_i = int(ref_idx)
                c2w_ref = self.pose_param_net(ref_idx_i, ref_idx_dev)
                ss_ref_raw = self.distortion_net.scale_shift(ref_idx_i, ref_idx_dev) if self.distortion_net is not None else None
                if self.detach_ref_img:
                    c2w_ref = c2w_ref.detach()
                    ss_ref_raw = ss_ref_raw.detach() if ss_ref_raw is not None else None
                ss_ref = affine(ss_ref_raw)
                if img_idx_i < (num_cams - 1):
                    d1_raw, d2_raw, ss1, ss2 = depth_raw, depth_ref_raw, ss_in, ss_ref
                    img1, img2 = img, ref_img
                    # ref_Rt @ inverse(world_mat) with inverse(world_mat) == c2w (model/training.py:329)
                    Rt_rel_12 = _ops.Mat4Mul.apply(_ops.Mat4Inverse.apply(c2w_ref), c2w)
                else:
                    d1_raw, d2_raw, ss1, ss2 = depth_ref_raw, depth_raw, ss_ref, ss_in
                    img1, img2 = ref_img, img
                    Rt_rel_12 = _ops.Mat4Mul.apply(world_mat[0], c2w_ref)                 # world_mat @ inverse(ref_Rt) (model/training.py:344)
                R_rel_12 = Rt_rel_12[:3, :3]
                t_rel_12 = Rt_rel_12[:3, 3]
                scale1 = ss1[0] if ss1 is not None else None

                ratio = self.pc_ratio
                h4, w4 = int(h_depth / ratio), int(w_depth / ratio)
                pc1 = _ops.Backproject.apply(d1_raw, ss1, camera_mat, h4, w4, nl)   # (P,3)
                pc2 = _ops.Backproject.apply(d2_raw, ss2, camera_mat, h4, w4, nl)

                if weights['rgb_s_weight'] != 0.0:
                    img1r = _ops.resize_bilinear(img1[0], h4, w4)
                    img2r = _ops.resize_bilinear(img2[0], h4, w4)
                    pc1_in = pc1.detach() if self.detach_rgbs_scale else pc1
                    rgb_pc1, rgb_pc1_proj, valid = _ops.Reproject.apply(img1r, img2r, pc1_in, R_rel_12, t_rel_12, camera_mat, nl)
                    kwargs['rgb_pc1'] = rgb_pc1.view(batch_size, h4, w4, 3)
                    kwargs['rgb_pc1_proj'] = rgb_pc1_proj.view(batch_size, h4, w4, 3)
                    kwargs['valid_points'] = valid.view(batch_size, h4, w4, 1)
                # X = (pc1 / scale1) @ R^T + t, Y = pc2 / scale1 in one launch (model/training.py:394-399)
                Xa, Ya = _ops.AlignClouds.apply(pc1, pc2, Rt_rel_12, scale1 if self.scale_pcs else None)
                kwargs['X'] = Xa.unsqueeze(0)
                kwargs['Y'] = Ya.unsqueeze(0)
                kwargs['sample_resolution'] = (h4, w4)
                kwargs['rt_12'] = Rt_rel_12.unsqueeze(0)
                if weights['t_cycle_weight'] != 0.0:
                    world_mat_gt = torch.inverse(pose_gt).unsqueeze(0)
                    ref_Rt_gt = torch.inverse(ref_pose_gt).unsqueeze(0)
                    if img_idx_i < (num_cams - 1):
                        kwargs['rt_12_gt'] = ref_Rt_gt @ torch.inverse(world_mat_gt)
                    else:
                        kwargs['rt_12_gt'] = world_mat_gt @ torch.inverse(ref_Rt_gt)
                if geom is not None:
                    # the two losses of this branch are evaluated here, on its stream (model/losses.py:111-112); the photometric
                    # term first: the main stream waits for the END of this branch, which then is the Chamfer search's own tail
                    if weights['rgb_s_weight'] != 0.0:
                        kwargs['rgb_s_loss_pre'] = self.loss.get_rgb_s_loss(kwargs['rgb_pc1'], kwargs['rgb_pc1_proj'], kwargs['valid_points'])
                    if weights['pc_weight'] != 0.0:
                        kwargs['pc_loss_pre'] = self.loss.get_pc_loss(kwargs['X'], kwargs['Y'])

        # Ray choice (model/training.py:277); ranks draw the SAME global batch.  'reference': the reference's own call,
        # torch.randperm(h*w, device)[:n] on the device generator, followed by its torch.rand(1, R, S) (model/rendering.py:190):
        # a seeded run visits the reference's rays and jitter (tests/test_gpu_step.py::test_seeded_draws_match_reference_calls).
        # 'feistel' (default): head of a keyed pseudo-random permutation, no sort of h*w keys (240 us at 375 x 1242); it
        # consumes the generator differently (six randint keys), so seeds are NOT reference-compatible in that mode.
        def draw():
            if self.ray_choice == 'reference':
                return torch.randperm(h * w, device=device)[:n_points]
            return _ops.ray_choice(h * w, min(n_points, h * w), device)
        if self.fixed_ray_idx is not None:
            ray_idx = self.fixed_ray_idx.to(device)
        else:
            ray_idx = draw()
            if data.get('img.dpt') is None:
                # sparse depth maps (dataloading `with_depth` + `sparsify_depth`, the V_KITTI configs): the reference draws again
                # until at least one chosen pixel has a depth (model/training.py:279-284), reading the mask on the host
                depth_mask_flat = torch.flatten(depth_mask)
                if self._graph_body:
                    # graph-safe form: a second draw, selected on the device when the first holds no valid pixel (probability
                    # (1 - valid fraction)^n_rays: 1e-22 for 4,096 rays on a 1-in-81 pattern); both empty -> flag, checked
                    # one step late with the NaN check (Loss.raise_if_nan)
                    alt = draw()
                    ok = depth_mask_flat[ray_idx].any()
                    self._no_valid_depth = (~(ok | depth_mask_flat[alt].any())).float()
                    ray_idx = torch.where(ok, ray_idx, alt)
                else:
                    while not depth_mask_flat[ray_idx].any():
                        print('No valid depths found, resampling...')
                        ray_idx = draw()
        n_global = ray_idx.numel()
        lo, hi = self.dp.shard(n_global)
        sharded = self.dp.world_size > 1
        if hi <= lo:
            raise ValueError("data parallelism over %d ranks needs at least that many rays per step (n_training_points = %d): rank %d "
                             "would render an empty shard" % (self.dp.world_size, n_global, self.dp.rank))
        shard = (lo, hi) if sharded else None

        if (h_depth, w_depth) != (h, w):
            # model/network.py:21-24 resamples the depth map to the image size (F.interpolate, mode 'area') before the gather
            depth_raw = torch.nn.functional.interpolate(depth_raw, (h, w), mode='area')
        # fused gather: pixels, rgb_gt and depth*scale+shift for the selected rays only.  Data parallel: stage 1 (33 B/ray)
        # covers the GLOBAL batch on every rank -- the validity mask of the whole batch is then known without a collective --
        # and only this rank's rays [lo, hi) are rendered (Renderer.nope_nerf `shard`).
        pixels, rgb_gt, depth_sel = _ops.GatherRays.apply(ray_idx.contiguous(), img[0], depth_raw[0, 0], ss_in, h, w)
        rgb_gt = (rgb_gt[lo:hi] if sharded else rgb_gt).unsqueeze(0)

        out_dict = None
        if render_model:
            renderer = self.model.renderer
            S = renderer.cfg['num_points'] - renderer.cfg['outside_steps']
            if self.fixed_noise is not None:
                noise = self.fixed_noise.to(device).reshape(n_global, S)
            elif not eval_mode:
                # the reference draws (1, R, S) inside the renderer (model/rendering.py:190)
                noise = torch.rand(1, n_global, S, device=device).reshape(n_global, S)
            else:
                noise = None
            self._last_draw = (ray_idx, noise)    # references only (parity tests read them)
            out_dict = renderer(pixels.unsqueeze(0), depth_sel.reshape(1, -1, 1), camera_mat, world_mat, scale_mat,
                                self.rendering_technique, add_noise=noise is not None, eval_=eval_mode, it=it, noise=noise,
                                lazy_mask=True, shard=shard)
            rendered_rgb = out_dict['rgb']
            rendered_depth = out_dict['depth_pred_dense']
            gt_depth = out_dict['depth_gt_dense']
            kwargs['depth_mask'] = out_dict['mask']
            if sharded:
                gt_all = out_dict['depth_gt_global']
                kwargs['dp_info'] = {'lo': lo, 'hi': hi, 'n_global': n_global, 'mask_global': out_dict['mask_global'],
                                     'depth_gt_global': gt_all.detach() if self.detach_gt_depth else gt_all}
        else:
            rendered_rgb = rendered_depth = gt_depth = None
            if sharded:
                kwargs['dp_info'] = {'lo': lo, 'hi': hi, 'n_global': n_global}

        if geom is not None:
            kwargs['join_pre'] = lambda: main_stream.wait_stream(geom)   # Loss.forward joins the branch where it first needs it
        if render_model and self.detach_gt_depth:
            gt_depth = gt_depth.detach()
        loss_dict = self.loss(rendered_rgb, rgb_gt, rendered_depth, gt_depth, **kwargs)
        if self.optimizer_focal:
            loss_dict['focalx'] = fxfy[0] / camera_mat_gt[0, 0, 0]
            loss_dict['focaly'] = fxfy[1] / camera_mat_gt[0, 1, 1]
        if ss_raw is not None:
            loss_dict['scale'] = ss_raw[0:1]
            loss_dict['shift'] = ss_raw[1:2]
        return loss_dict

    def render_visdata(self, data, resolution, it, out_render_path):
        """Full-frame preview (model/training.py:103-165, called by train.py:239 every `visualize_every` iterations): renders frame
        `img.idx` at `resolution`, writes `<idx>_depth.png` (min-max normalised), `<idx>_img.png` and -- with `vis_geo`, the
        default -- `<idx>_geo.png` (phong-shaded sphere-traced surface) into `out_render_path`, and returns the last image
        written as uint8 (h,w,3), like the reference."""
        (img, depth_input, camera_mat, scale_mat, img_idx, *other) = self.process_data_dict(data)
        h, w = resolution
        idx = int(img_idx)
        c2w = self.pose_param_net(idx)
        world_mat = _ops.Mat4Inverse.apply(c2w.detach()).unsqueeze(0)
        _, pixels = _ops.pixel_grid(h, w, self.device)
        with torch.no_grad():
            out = self.model.renderer(pixels.unsqueeze(0), torch.ones(1, h * w, 1, device=self.device), camera_mat, world_mat,
                                      scale_mat, self.rendering_technique, add_noise=False, eval_=True, it=it)
        img_out = (out['rgb'].view(h, w, 3) * 255).to(torch.uint8).cpu().numpy()
        depth = out['depth_pred'].view(h, w).cpu().numpy()
        depth_u8 = np.clip(255.0 / depth.max() * (depth - depth.min()), 0, 255).astype(np.uint8)
        dp_rank = self.dp.rank if self.dp is not None else 0
        if out_render_path is not None and dp_rank == 0:
            _imwrite(os.path.join(out_render_path, '%04d_depth.png' % idx), depth_u8)
            _imwrite(os.path.join(out_render_path, '%04d_img.png' % idx), img_out)
        if self.vis_geo:
            with torch.no_grad():
                geo = self.model(pixels.unsqueeze(0), None, camera_mat, world_mat, scale_mat, 'phong_renderer', add_noise=False,
                                 eval_mode=True, it=it, depth_img=depth_input, img_size=(h, w))['rgb']
            img_out = (geo.view(h, w, 3) * 255).to(torch.uint8).cpu().numpy()
            if out_render_path is not None and dp_rank == 0:
                _imwrite(os.path.join(out_render_path, '%04d_geo.png' % idx), img_out)
        return img_out
