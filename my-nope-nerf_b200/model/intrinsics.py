"""LearnFocal -- same parametrisation and state_dict keys (`fx`, `fy`) as the reference's model/intrinsics.py.
Out of the kernel scope (learn_focal is False in every shipped config): it is two
scalars, kept as plain PyTorch.  The ray-generation kernel returns d(camera_mat), so gradients reach it."""
import numpy as np
import torch
import torch.nn as nn


def _coef(v, order):
    if order == 2:
        return torch.tensor(np.sqrt(v), requires_grad=False).float()
    if order == 1:
        return torch.tensor(v, requires_grad=False).float()
    raise ValueError("Focal init order needs to be 1 or 2")  # the reference calls exit() here


class LearnFocal(nn.Module):
    def __init__(self, req_grad, fx_only, order=2, init_focal=None):
        super().__init__()
        self.fx_only = fx_only
        self.order = order
        if init_focal is None:
            fx = fy = torch.tensor(1.0, dtype=torch.float32)
        elif isinstance(init_focal, list) and not fx_only:
            fx, fy = _coef(init_focal[0], order), _coef(init_focal[1], order)
        else:
            fx = fy = _coef(init_focal, order)
        self.fx = nn.Parameter(fx.clone(), requires_grad=req_grad)
        if not fx_only:
            self.fy = nn.Parameter(fy.clone(), requires_grad=req_grad)

    def forward(self, i=None):
        fx = self.fx
        fy = self.fx if self.fx_only else self.fy
        if self.order == 2:
            return torch.stack([fx ** 2, fy ** 2])
        return torch.stack([fx, fy])
