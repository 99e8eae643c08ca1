"""Learn_Distortion -- drop-in for the reference's model/distortions.py (state_dict keys `global_scales`,
`global_shifts`).  The scale clamp and the last-camera pin happen inside the kernel, so there is no host sync
(the reference's `if scale<0.01` reads the value back)."""
import torch
import torch.nn as nn

from . import _ops


class Learn_Distortion(nn.Module):
    def __init__(self, num_cams, learn_scale, learn_shift, cfg):
        super().__init__()
        self.global_scales = nn.Parameter(torch.ones(size=(num_cams, 1), dtype=torch.float32), requires_grad=learn_scale)
        self.global_shifts = nn.Parameter(torch.zeros(size=(num_cams, 1), dtype=torch.float32), requires_grad=learn_shift)
        self.fix_scaleN = cfg['distortion']['fix_scaleN']
        self.num_cams = num_cams

    def scale_shift(self, cam_id, cam_id_dev=None):
        """Device tensor [scale, shift] (what the fused gather / back-projection kernels consume).  With `cam_id_dev`
        (int64 device tensor holding the same index) the row is gathered on the device; only `cam_id == num_cams - 1`
        (the fix_scaleN pin) is then decided on the host."""
        cam_id = int(cam_id)
        if cam_id_dev is not None:
            pin = bool(self.fix_scaleN) and cam_id == self.num_cams - 1
            return _ops.Distortion.apply(self.global_scales.index_select(0, cam_id_dev),
                                         self.global_shifts.index_select(0, cam_id_dev), 0, pin)
        return _ops.Distortion.apply(self.global_scales, self.global_shifts, cam_id, self.fix_scaleN)

    def forward(self, cam_id):
        ss = self.scale_shift(cam_id)
        return ss[0:1], ss[1:2]
