"""nope_nerf -- drop-in for the reference's model/network.py (the module `get_model` builds; its parameters are what
train.py hands to Adam and what model.pt stores under `renderer.model.*`)."""
import torch.nn as nn
import torch.nn.functional as F


class nope_nerf(nn.Module):
    def __init__(self, cfg, renderer, depth_estimator=None, device=None, **kwargs):
        super().__init__()
        self.renderer = renderer.to(device)
        self.depth_estimator = depth_estimator.to(device) if depth_estimator is not None else None
        self.device = device

    def forward(self, p, ray_idx, camera_mat, world_mat, scale_mat, rendering_technique, it=0, eval_mode=False,
                depth_img=None, add_noise=True, img_size=None, noise=None):
        if rendering_technique == 'nope_nerf':
            if tuple(depth_img.shape[-2:]) != tuple(img_size):
                # only differs from the identity when the depth map is stored at another resolution
                depth_img = F.interpolate(depth_img, img_size, mode='area')
            depth = depth_img.reshape(1, -1, 1)[:, ray_idx]  # model/network.py:21-24
        else:
            depth = None
        return self.renderer(p, depth, camera_mat, world_mat, scale_mat, rendering_technique, eval_=eval_mode, it=it,
                             add_noise=add_noise, noise=noise)
