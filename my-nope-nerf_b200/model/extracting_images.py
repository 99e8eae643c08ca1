"""Extract_Images -- the full-frame novel-view render loop of the reference's model/extracting_images.py:54-76.
Pixel rows can be sharded across ranks (parallel.DataParallelContext); `generate_images` returns and writes what the
reference's does (img / depth / disp / geo), the MP4 assembly stays in vis/render.py."""
import os

import numpy as np
import torch

from . import _ops, parallel


class Extract_Images(object):
    def __init__(self, renderer, cfg, use_learnt_poses=True, use_learnt_focal=True, device=None, render_type=None, dp=None):
        self.points_batch_size = 100000
        self.renderer = renderer
        self.resolution = cfg['extract_images']['resolution']
        self.device = device
        self.use_learnt_poses = use_learnt_poses
        self.use_learnt_focal = use_learnt_focal
        self.render_type = render_type
        self.dp = dp or parallel.DataParallelContext.from_env()

    def process_data_dict(self, data):
        device = self.device
        return (data.get('img.camera_mat').to(device), data.get('img.scale_mat').to(device), data.get('img.idx'))

    def render_frame(self, camera_mat, world_mat, scale_mat, h, w, it=0, gather=True):
        """-> rgb (h,w,3) fp32, depth (h,w) fp32 on the device.  With world_size > 1 each rank renders a contiguous pixel
        shard; `gather` assembles the full frame on every rank."""
        n = h * w
        lo, hi = self.dp.shard(n)
        _, pixels = _ops.pixel_grid(h, w, self.device)
        pixels = pixels[lo:hi]
        rgb_pred, depth_pred = [], []
        with torch.no_grad():
            for a in range(0, pixels.shape[0], self.points_batch_size):
                p = pixels[a:a + self.points_batch_size].unsqueeze(0)
                d = torch.ones(1, p.shape[1], 1, device=self.device)  # depth 1: avoids masking (reference :60-62)
                out = self.renderer(p, d, camera_mat, world_mat, scale_mat, self.render_type, eval_=True, it=it,
                                    add_noise=False)
                rgb_pred.append(out['rgb'][0])
                depth_pred.append(out['depth_pred'])
        rgb = torch.cat(rgb_pred, dim=0)
        depth = torch.cat(depth_pred, dim=0)
        if gather and self.dp.world_size > 1:
            rgb = self.dp.gather_rows(rgb, n)
            depth = self.dp.gather_rows(depth, n)
            return rgb.view(h, w, 3), depth.view(h, w)
        if self.dp.world_size > 1:
            return rgb, depth
        return rgb.view(h, w, 3), depth.view(h, w)

    def generate_images(self, data, render_dir, c2ws, fxfy, it, output_geo):
        """Reference signature and outputs (model/extracting_images.py:40-131): renders frame `img.idx`, writes
        depth_out/<idx>.npy and the img / depth / disparity (/ geo) PNGs under `render_dir`, and returns
        {'img': uint8 (h,w,3), 'depth': uint8 (h,w) min-max normalised, 'geo': uint8 (h,w,3) or None, 'disp': uint8 (h,w,3)
        inferno-mapped 1/depth} -- the four arrays vis/render.py:106-109 stacks into videos."""
        self.renderer.eval()
        (camera_mat, scale_mat, img_idx) = self.process_data_dict(data)
        img_idx = int(img_idx)
        world_mat = _ops.Mat4Inverse.apply(c2ws[img_idx].to(self.device)).unsqueeze(0)
        if self.use_learnt_focal and fxfy is not None:
            camera_mat = torch.tensor([[[float(fxfy[0]), 0, 0, 0], [0, -float(fxfy[1]), 0, 0], [0, 0, -1, 0], [0, 0, 0, 1]]],
                                      device=self.device)
        h, w = self.resolution
        rgb, depth = self.render_frame(camera_mat, world_mat, scale_mat, h, w, it=it)
        img_out = (rgb * 255).to(torch.uint8).cpu().numpy()
        depth_raw = depth.cpu().numpy()

        geo_out = None
        if output_geo:
            # model/extracting_images.py:79-98: the phong-shaded sphere-traced surface, 1,024 pixels per call
            _, pixels = _ops.pixel_grid(h, w, self.device)
            with torch.no_grad():
                parts = [self.renderer(pixels[a:a + 1024].unsqueeze(0), None, camera_mat, world_mat, scale_mat, 'phong_renderer',
                                       eval_=True, it=it, add_noise=False)['rgb'] for a in range(0, h * w, 1024)]
            geo_out = (torch.cat(parts, dim=1).reshape(h, w, 3) * 255).to(torch.uint8).cpu().numpy()

        disp = 1 / depth_raw
        depth_out = np.clip(255.0 / depth_raw.max() * (depth_raw - depth_raw.min()), 0, 255).astype(np.uint8)
        disp_u8 = np.clip(255.0 / disp.max() * (disp - disp.min()), 0, 255).astype(np.uint8)
        disp_out = _inferno(disp_u8)
        if render_dir is not None and self.dp.rank == 0:
            name = str(img_idx).zfill(4) + '.png'
            dirs = {k: os.path.join(render_dir, k) for k in ('img_out', 'depth_out', 'disp_out', 'geo_out')}
            for k, d in dirs.items():
                if k != 'geo_out' or output_geo:
                    os.makedirs(d, exist_ok=True)
            np.save(os.path.join(dirs['depth_out'], '{}.npy'.format(img_idx)), depth_raw)
            _imwrite(os.path.join(dirs['img_out'], name), img_out)
            _imwrite(os.path.join(dirs['depth_out'], name), depth_out)
            _imwrite(os.path.join(dirs['disp_out'], name), disp_out)
            if output_geo:
                _imwrite(os.path.join(dirs['geo_out'], name), geo_out)
        return {'img': img_out, 'depth': depth_out, 'geo': geo_out, 'disp': disp_out}


def _inferno(gray_u8):
    """cv2.applyColorMap(x, COLORMAP_INFERNO)[:, :, ::-1] (model/extracting_images.py:120): uint8 (h,w) -> RGB uint8 (h,w,3)."""
    import cv2
    return np.ascontiguousarray(cv2.applyColorMap(gray_u8, cv2.COLORMAP_INFERNO)[:, :, ::-1])


def _imwrite(path, arr):
    """PNG output: imageio when it is installed (what the reference calls), else OpenCV (RGB -> BGR)."""
    try:
        import imageio
        if hasattr(imageio, 'imwrite'):
            imageio.imwrite(path, arr)
            return
    except ImportError:
        pass
    import cv2
    cv2.imwrite(path, arr[:, :, ::-1] if arr.ndim == 3 else arr)
