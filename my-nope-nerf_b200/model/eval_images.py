"""Eval_Images -- the per-frame evaluation loop body of the reference's model/eval_images.py (called by evaluation/eval.py:179).

The frame is rendered by the fused kernels (Extract_Images.render_frame); what follows is the reference's host-side bookkeeping
with the same return keys, file names and array conventions: colour MSE / PSNR / SSIM / LPIPS (the LPIPS network is the callable
evaluation/eval.py builds and passes in), the rendered depth rescaled by `sc` and resized (nearest-exact) to the ground-truth
depth map, the validity masks and the confusion matrix of `min_depth <= depth <= max_depth`, and the PNGs under `render_dir`."""
import logging
import os

import numpy as np
import torch
import torch.nn.functional as F

from . import _ops
from .common import mse2psnr
from .extracting_images import Extract_Images, _imwrite

logger_py = logging.getLogger(__name__)


def ssim(img1, img2, use_padding=True, window_size=11, size_average=True):
    """Mean SSIM over channels with an 11 x 11 Gaussian window of sigma 1.5 (third_party/pytorch_ssim/__init__.py:8-92, the
    function model/eval_images.py:97 calls).  img1, img2: (B,C,H,W) in [0,1]."""
    channel = img1.shape[1]
    x = torch.arange(window_size, dtype=torch.float32) - window_size // 2
    g = torch.exp(-x * x / (2 * 1.5 ** 2))
    g = (g / g.sum()).unsqueeze(1)
    window = (g @ g.t()).expand(channel, 1, window_size, window_size).contiguous().to(img1)
    pad = window_size // 2 if use_padding else 0
    blur = lambda t: F.conv2d(t, window, padding=pad, groups=channel)
    mu1, mu2 = blur(img1), blur(img2)
    s11, s22, s12 = blur(img1 * img1) - mu1 * mu1, blur(img2 * img2) - mu2 * mu2, blur(img1 * img2) - mu1 * mu2
    c1, c2 = 0.01 ** 2, 0.03 ** 2
    smap = ((2 * mu1 * mu2 + c1) * (2 * s12 + c2)) / ((mu1 * mu1 + mu2 * mu2 + c1) * (s11 + s22 + c2))
    return smap.mean() if size_average else smap.mean(dim=(1, 2, 3))


def _minmax_u8(a):
    return np.clip(255.0 / a.max() * (a - a.min()), 0, 255).astype(np.uint8)


class Eval_Images(Extract_Images):
    def __init__(self, renderer, cfg, points_batch_size=100000, use_learnt_poses=True, use_learnt_focal=True, device=None,
                 render_type=None, c2ws=None, img_list=None):
        super().__init__(renderer, cfg, use_learnt_poses=use_learnt_poses, use_learnt_focal=use_learnt_focal, device=device,
                         render_type=render_type)
        self.points_batch_size = points_batch_size
        self.c2ws = c2ws
        self.img_list = img_list

    def process_data_dict(self, data):
        device = self.device
        img = data.get('img').to(device)
        batch_size, _, h, w = img.shape
        depth_img = data.get('img.depth', torch.ones(batch_size, h, w))
        return (img, depth_img, data.get('img.camera_mat').to(device), data.get('img.scale_mat').to(device), data.get('img.idx'),
                data.get('img.gt_depths'))

    def eval_images(self, data, render_dir, fxfy, lpips_vgg_fn, logger, min_depth=0.1, max_depth=20, it=0, sc=1, show_errors=False):
        import cv2
        self.renderer.eval()
        (img_gt, depth, camera_mat, scale_mat, img_idx, depth_gt) = self.process_data_dict(data)
        img_idx = int(img_idx)
        img_gt = img_gt.squeeze(0).permute(1, 2, 0)
        if depth_gt is not None:
            depth_gt = depth_gt.squeeze(0).cpu().numpy()
        else:
            print('No GT depths available, using input depths')
            depth_gt = depth.squeeze(0).cpu().numpy()
        world_mat = _ops.Mat4Inverse.apply(self.c2ws[img_idx].to(self.device)).unsqueeze(0)
        if self.use_learnt_focal:
            camera_mat = torch.tensor([[[float(fxfy[0]), 0, 0, 0], [0, -float(fxfy[1]), 0, 0], [0, 0, -1, 0], [0, 0, 0, 1]]],
                                      device=self.device)
        h, w = self.resolution
        img_out, depth_dev = self.render_frame(camera_mat, world_mat, scale_mat, h, w, it=it)
        depth_out = depth_dev.cpu().numpy()

        mse = F.mse_loss(img_out, img_gt).item()
        psnr = mse2psnr(mse)
        a, b = img_out.permute(2, 0, 1).unsqueeze(0), img_gt.permute(2, 0, 1).unsqueeze(0)
        ssim_val = ssim(a, b).item()
        lpips_loss = lpips_vgg_fn(a.contiguous(), b.contiguous(), normalize=True).item()
        print('{0:4d} img: PSNR: {1:.2f}, SSIM: {2:.2f},  LPIPS {3:.2f}'.format(img_idx, psnr, ssim_val, lpips_loss))

        gt_height, gt_width = depth_gt.shape[:2]
        depth_out = depth_out * sc
        depth_out = cv2.resize(depth_out, (gt_width, gt_height), interpolation=cv2.INTER_NEAREST_EXACT)

        dirs = {k: os.path.join(render_dir, k) for k in ('img_out', 'depth_out', 'img_gt_out', 'depth_gt_out', 'disp_out', 'disp_gt_out',
                                                          'depth_mask')}
        for d in dirs.values():
            os.makedirs(d, exist_ok=True)
        depth_img, depth_img_gt = _minmax_u8(depth_out), _minmax_u8(depth_gt)
        img_out = (img_out.cpu().numpy() * 255).astype(np.uint8)
        img_gt = (img_gt.cpu().numpy() * 255).astype(np.uint8)
        disp_img_gt = cv2.applyColorMap(_minmax_u8(1 / depth_gt), cv2.COLORMAP_INFERNO)
        disp_img = cv2.applyColorMap(_minmax_u8(1 / depth_out), cv2.COLORMAP_INFERNO)
        name = str(img_idx).zfill(4)
        _imwrite(os.path.join(dirs['img_out'], name + '.png'), img_out)
        _imwrite(os.path.join(dirs['depth_out'], name + '.png'), depth_img)
        _imwrite(os.path.join(dirs['img_gt_out'], name + '.png'), img_gt)
        _imwrite(os.path.join(dirs['depth_gt_out'], name + '.png'), depth_img_gt)
        cv2.imwrite(os.path.join(dirs['disp_gt_out'], name + '.png'), disp_img_gt)      # BGR arrays through OpenCV, as the reference
        cv2.imwrite(os.path.join(dirs['disp_out'], name + '.png'), disp_img)

        mask_rendered = (depth_out >= min_depth) & (depth_out <= max_depth)
        mask_gt = (depth_gt >= min_depth) & (depth_gt <= max_depth)
        mask = mask_rendered & mask_gt
        tp = (mask_rendered & mask_gt).reshape(1, -1)
        tn = (~mask_rendered & ~mask_gt).reshape(1, -1)
        fp = (mask_rendered & ~mask_gt).reshape(1, -1)
        fn = (~mask_rendered & mask_gt).reshape(1, -1)
        num_pixels = np.arange(depth_gt.shape[0] * depth_gt.shape[1]).reshape(1, -1)
        if show_errors:
            self._plot_errors(os.path.join(render_dir, name + '_conf.png'), num_pixels, (depth_gt - depth_out).reshape(1, -1), tp, tn, fp, fn)

        def masked_rgb(gray):
            """pixels outside the evaluated range in green, the others keep their grey value"""
            rb, g = gray.copy(), gray.copy()
            rb[mask == 0] = 0
            g[mask == 0] = 255
            return np.stack((rb, g, rb), axis=-1)
        _imwrite(os.path.join(dirs['depth_mask'], name + '_mask_rendered.png'), (255 * mask_rendered).astype(np.uint8))
        _imwrite(os.path.join(dirs['depth_mask'], name + '_mask_gt.png'), (255 * mask_gt).astype(np.uint8))
        _imwrite(os.path.join(dirs['depth_mask'], name + '_mask_combined.png'), (255 * mask).astype(np.uint8))
        _imwrite(os.path.join(dirs['depth_mask'], name + '_gt.png'), masked_rgb(depth_img_gt))
        _imwrite(os.path.join(dirs['depth_mask'], name + '.png'), masked_rgb(depth_img))

        return {'img': img_out, 'mse': mse, 'psnr': psnr, 'ssim': ssim_val, 'lpips': lpips_loss,
                'depth_pred': depth_out[mask], 'depth_gt': depth_gt[mask],
                'conf_mat': np.array([[np.sum(tp), np.sum(fn)], [np.sum(fp), np.sum(tn)]]) / num_pixels[0, -1]}

    @staticmethod
    def _plot_errors(path, num_pixels, x, tp, tn, fp, fn):
        """The scatter plot of model/eval_images.py:159-172 (first frame only); skipped with a warning without matplotlib."""
        try:
            import matplotlib.pyplot as plt
        except ImportError:
            logger_py.warning('matplotlib is not installed: %s not written' % path)
            return
        plt.figure()
        plt.xlim(num_pixels[0, 0], num_pixels[0, -1])
        for sel, colour in ((tp, 'r'), (tn, 'g'), (fp, 'b'), (fn, 'k')):
            plt.scatter(num_pixels[sel], x[sel], 1, colour)
        plt.legend(['True Positive', 'True Negative', 'False Positive', 'False Negative'])
        plt.xlabel('Pixel Index')
        plt.ylabel('GT Depth - Predicted Depth (m)')
        plt.title('Classification of Depth Errors')
        plt.savefig(path)
