"""LearnPose -- drop-in for the reference's model/poses.py (same constructor, forward, get_t, state_dict keys
`init_c2w` (only when given), `r`, `t`).  forward() is one launch of the SE(3) exp-map kernel."""
import torch
import torch.nn as nn

from . import _ops


class LearnPose(nn.Module):
    def __init__(self, num_cams, learn_R, learn_t, cfg, init_c2w=None):
        """
        :param num_cams: number of cameras
        :param learn_R / learn_t: whether the axis-angle / translation parameters take gradients
        :param cfg: config (unused, kept for signature parity; reference model/poses.py:7)
        :param init_c2w: (N, 4, 4) frozen initial poses or None
        """
        super().__init__()
        self.num_cams = num_cams
        self.init_c2w = None
        if init_c2w is not None:
            self.init_c2w = nn.Parameter(init_c2w, requires_grad=False)
        self.r = nn.Parameter(torch.zeros(size=(num_cams, 3), dtype=torch.float32), requires_grad=learn_R)
        self.t = nn.Parameter(torch.zeros(size=(num_cams, 3), dtype=torch.float32), requires_grad=learn_t)

    def forward(self, cam_id, cam_id_dev=None):
        """`cam_id_dev` (optional, int64 device tensor of one element holding the same index): the rows are then gathered
        on the device, so that nothing in the launch sequence depends on the host value of the index -- what a CUDA graph
        of the training step needs (Trainer.train_step_graphed)."""
        cam_id = int(cam_id)
        if cam_id < 0 or cam_id >= self.num_cams:
            raise IndexError("cam_id %d out of range for %d cameras" % (cam_id, self.num_cams))
        if cam_id_dev is not None:
            init = self.init_c2w.index_select(0, cam_id_dev) if self.init_c2w is not None else None
            return _ops.PoseC2W.apply(self.r.index_select(0, cam_id_dev), self.t.index_select(0, cam_id_dev), init, 0)
        return _ops.PoseC2W.apply(self.r, self.t, self.init_c2w, cam_id)

    def get_t(self):
        return self.t
