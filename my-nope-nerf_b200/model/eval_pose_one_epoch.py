"""Trainer_pose -- drop-in for the reference's model/eval_pose_one_epoch.py: freezes the MLP and fits test poses on
the rgb MSE.  A second caller of the same boundary."""
import torch

from . import _ops
from .losses import Loss_Eval


class Trainer_pose(object):
    def __init__(self, model, cfg, device=None, optimizer_pose=None, pose_param_net=None, focal_net=None, **kwargs):
        self.model = model
        self.device = device
        self.optimizer_pose = optimizer_pose
        self.pose_param_net = pose_param_net
        self.focal_net = focal_net
        self.n_points = cfg['n_points']
        self.rendering_technique = cfg['type']
        self.loss = Loss_Eval()

    def train_step(self, data, it=100000):
        self.model.eval()
        self.pose_param_net.train()
        self.optimizer_pose.zero_grad()
        if self.focal_net is not None:
            self.focal_net.eval()
        loss_dict = self.compute_loss(data, it=it)
        loss_dict['loss'].backward()
        self.optimizer_pose.step()
        return loss_dict

    def process_data_dict(self, data):
        device = self.device
        img = data.get('img').to(device)
        img_idx = data.get('img.idx').to(device)
        batch_size, _, h, w = img.shape
        depth_img = data.get('img.depth', torch.ones(batch_size, h, w)).unsqueeze(1).to(device)
        camera_mat = data.get('img.camera_mat').to(device)
        scale_mat = data.get('img.scale_mat').to(device)
        return (img, depth_img, camera_mat, scale_mat, img_idx)

    def compute_loss(self, data, eval_mode=False, it=100000):
        (img, depth_img, camera_mat, scale_mat, img_idx) = self.process_data_dict(data)
        device = self.device
        _, _, h, w = img.shape
        c2w = self.pose_param_net(int(img_idx))
        world_mat = _ops.Mat4Inverse.apply(c2w).unsqueeze(0)
        if self.focal_net is not None:
            fxfy = self.focal_net(0)
            pad = torch.zeros(4, device=device)
            one = torch.ones(1, device=device)
            camera_mat = torch.cat([fxfy[0:1], pad, -fxfy[1:2], pad, -one, pad, one]).view(1, 4, 4)
        ray_idx = torch.randperm(h * w, device=device)[:self.n_points]
        pixels, rgb_gt, depth = _ops.GatherRays.apply(ray_idx, img[0], depth_img[0, 0], None, h, w)
        out_dict = self.model.renderer(pixels.unsqueeze(0), depth.reshape(1, -1, 1), camera_mat, world_mat, scale_mat,
                                       self.rendering_technique, it=it, eval_=True, add_noise=False)
        return self.loss(out_dict['rgb'], rgb_gt.unsqueeze(0))
