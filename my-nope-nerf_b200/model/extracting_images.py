"""Extract_Images -- the full-frame novel-view render loop of the reference's model/extracting_images.py:54-76.
Pixel rows can be sharded across ranks (parallel.DataParallelContext); file output (PNG/NPY/MP4) stays with the
caller -- it is host I/O outside the hot path."""
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
        """Reference signature (model/extracting_images.py:40).  Returns {'img': uint8 (h,w,3), 'depth': fp32 (h,w)} as
        numpy arrays; writing them to `render_dir` is the caller's job."""
        self.renderer.eval()
        (camera_mat, scale_mat, img_idx) = self.process_data_dict(data)
        img_idx = int(img_idx)
        world_mat = _ops.Mat4Inverse.apply(c2ws[img_idx].to(self.device)).unsqueeze(0)
        if self.use_learnt_focal and fxfy is not None:
            camera_mat = torch.tensor([[[float(fxfy[0]), 0, 0, 0], [0, -float(fxfy[1]), 0, 0], [0, 0, -1, 0], [0, 0, 0, 1]]],
                                      device=self.device)
        h, w = self.resolution
        rgb, depth = self.render_frame(camera_mat, world_mat, scale_mat, h, w, it=it)
        return {'img': (rgb * 255).to(torch.uint8).cpu().numpy(), 'depth': depth.cpu().numpy()}
