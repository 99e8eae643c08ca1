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
