"""ResidentScene -- every frame of a sequence kept on the device.

The reference moves one frame pair from a pinned-memory DataLoader to the GPU every step (dataloading/dataset.py:281-364,
model/training.py:172-181: 17 MB per step at 375 x 1242).  A sequence is small (20 frames x 375 x 1242 x (3 rgb + 1 depth +
1 gt depth) fp32 + masks = 190 MB), so it can live in HBM for the whole run; a step's batch is then a dict of VIEWS into the
resident tensors, built with the reference's batch schema and its reference-frame rule
(dataloading/dataset.py:295-301: `ref = idx - 1` for the last frame, else `idx + randint(1, min(random_ref, N - idx - 1))`).
`Trainer.train_step` / `train_step_graphed` take that dict unchanged; only the two frame indices cross the host boundary.
"""
import random

import torch

# per-frame entries of the batch schema, stacked along dim 0; batch dim of 1 restored by `batch()`
_PER_FRAME = ('img', 'img.dpt', 'img.depth', 'img.depth_mask', 'img.gt_depths', 'img.pose_gt')
_SHARED = ('img.camera_mat', 'img.scale_mat')
_REF_OF = {'img': 'img.ref_imgs', 'img.dpt': 'img.ref_dpts', 'img.depth': 'img.ref_depths', 'img.pose_gt': 'img.ref_pose_gt'}


class ResidentScene(object):
    def __init__(self, frames, device, random_ref=1):
        """frames: iterable of batch dicts as the reference's DataLoader yields them (batch size 1, any order, each with
        `img.idx`); reference-frame entries in them are ignored.  Everything is uploaded once."""
        frames = sorted(frames, key=lambda b: int(b['img.idx']))
        idxs = [int(b['img.idx']) for b in frames]
        if idxs != list(range(len(frames))):
            raise ValueError("ResidentScene needs every frame 0..N-1 exactly once (got %r)" % (idxs,))
        self.device = device
        self.random_ref = random_ref
        self.n_frames = len(frames)
        self.data = {}
        for k in _PER_FRAME:
            if k in frames[0] and frames[0][k] is not None:
                rows = [b[k].reshape(1, 4, 4) if k == 'img.pose_gt' else b[k] for b in frames]   # leading batch dim of 1
                self.data[k] = torch.cat(rows, 0).to(device).contiguous()
        for k in _SHARED:
            self.data[k] = frames[0][k].to(device).contiguous()
        self._idx_host = [torch.tensor([i]) for i in range(self.n_frames)]

    def nbytes(self):
        return sum(v.numel() * v.element_size() for v in self.data.values())

    def ref_index(self, idx, rng=random):
        """dataloading/dataset.py:295-301."""
        n = self.n_frames
        if idx == n - 1:
            return idx - 1
        return idx + rng.randint(1, min(self.random_ref, n - idx - 1))

    def batch(self, idx, ref_idx=None, with_ref=True):
        """The batch dict of frame `idx` (and its reference frame): views, no copies, no host synchronisation."""
        if with_ref and ref_idx is None:
            ref_idx = self.ref_index(idx)
        b = {'img.idx': self._idx_host[idx]}
        for k, v in self.data.items():
            if k in _SHARED:
                b[k] = v
            else:
                b[k] = v[idx:idx + 1]
        if with_ref and ref_idx is not None:
            b['img.ref_idxs'] = self._idx_host[ref_idx]
            for k, rk in _REF_OF.items():
                if k in self.data:
                    b[rk] = self.data[k][ref_idx:ref_idx + 1]
        return b


class BatchUploader(object):
    """Host -> device pipeline for batches that do not live on the device: `depth` pre-allocated sets of staging tensors, filled
    by asynchronous copies from the (pinned) host batch on a copy stream while the previous step computes.  No per-step device
    allocation (a growing caching-allocator pool costs a synchronising cudaMalloc per new block: measured 5 ms per step over the
    first ~20 steps when every step staged its batch in fresh tensors).

        up = BatchUploader(trainer, first_host_batch)
        up.submit(host_batches[0])
        for it in range(n):
            batch = up.next()                       # compute stream now waits for the upload of step `it`
            if it + 1 < n: up.submit(host_batches[it + 1])
            trainer.train_step_graphed(batch, ...)
            up.done()                               # the slot may be overwritten once this step has consumed it
    """

    def __init__(self, trainer, example, depth=2):
        self.keys = [k for k, v in example.items() if k in trainer.STEP_KEYS and torch.is_tensor(v) and k not in ('img.idx', 'img.ref_idxs')]
        self.device = trainer.device
        self.stream = torch.cuda.Stream(device=self.device)
        self.slots = [{k: torch.empty_like(example[k], device=self.device) for k in self.keys} for _ in range(depth)]
        self.uploaded = [None] * depth      # event: H2D copies of the slot are complete
        self.consumed = [None] * depth      # event: the step that read the slot has been issued and (when reached) finished
        self.meta = [None] * depth
        self.head = self.tail = 0

    def nbytes(self):
        return sum(v.numel() * v.element_size() for v in self.slots[0].values())

    def submit(self, host_batch):
        s = self.head % len(self.slots)
        self.head += 1
        with torch.cuda.stream(self.stream):
            if self.consumed[s] is not None:
                self.stream.wait_event(self.consumed[s])
            for k in self.keys:
                self.slots[s][k].copy_(host_batch[k], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self.uploaded[s] = ev
        self.meta[s] = {k: v for k, v in host_batch.items() if k in ('img.idx', 'img.ref_idxs')}

    def next(self):
        s = self.tail % len(self.slots)
        torch.cuda.current_stream().wait_event(self.uploaded[s])
        b = dict(self.slots[s])
        b.update(self.meta[s])
        return b

    def done(self):
        s = self.tail % len(self.slots)
        self.tail += 1
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self.consumed[s] = ev
