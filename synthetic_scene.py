"""Seeded synthetic scenes of the shapes BASELINE.json names.

Everything is generated on the CPU with fixed `torch.Generator`s so the CPU oracle and the
GPU path see identical bits.  Used by tests/, bench.py and oracle/make_golden.py; it has no
dependency on the reference, the oracle or the CUDA extension.
"""
import math

import torch

# name -> (frames, H, W, rays/step)
SHAPES = {
    "cfg1": (8, 120, 160, 1024),       # configs/default.yaml plumbing case
    "cfg2": (20, 375, 1242, 4096),     # Virtual-KITTI-shaped, 1 GPU
    "cfg3": (20, 375, 1242, 65536),    # same scene, 65,536-ray global batch
    "tiny": (4, 24, 32, 64),           # golden-fixture size
}

WEIGHT_NAMES = ["rgb_weight", "depth_weight", "pc_weight", "rgb_s_weight", "depth_consistency_weight",
                "weight_dist_2nd_loss", "weight_dist_1st_loss", "t_cycle_weight"]


def default_cfg():
    """The subset of configs/default.yaml the hot path reads (values copied from that file)."""
    return {
        "model": {"hidden_dim": 256, "pos_enc_levels": 10, "dir_enc_levels": 4,
                  "occ_activation": "softplus", "num_layers": 8},
        "rendering": {"type": "nope_nerf", "n_max_network_queries": 64000, "white_background": False,
                      "radius": 4.0, "num_points": 128, "depth_range": [0.01, 10], "dist_alpha": False,
                      "use_ray_dir": True, "normalise_ray": True, "normal_loss": False,
                      "sample_option": "uniform", "outside_steps": 0},
        "depth": {"type": "None"},
        "distortion": {"learn_distortion": True, "fix_scaleN": True, "learn_scale": True, "learn_shift": True},
        "pose": {"learn_pose": True, "learn_R": True, "learn_t": True, "init_pose": False, "learn_focal": False},
        "training": {
            "type": "nope_nerf", "n_training_points": 1024, "rgb_weight": [1.0, 1.0],
            "depth_weight": [0.04, 0.0], "weight_dist_2nd_loss": [0.0, 0.0], "weight_dist_1st_loss": [0.0, 0.0],
            "pc_weight": [1.0, 0.0], "rgb_s_weight": [1.0, 0.0], "depth_consistency_weight": [0.0, 0.0],
            "t_cycle_weight": [0.0, 0.0], "rgb_loss_type": "l1", "depth_loss_type": "l1",
            "with_auto_mask": False, "vis_geo": False, "with_ssim": False, "detach_gt_depth": False,
            "match_method": "dense", "pc_ratio": 4, "shift_first": False, "detach_ref_img": True,
            "scheduling_start": 10000, "scale_pcs": True, "detach_rgbs_scale": False,
            "vis_reprojection_every": 10 ** 9, "nearest_limit": 0.01, "annealing_epochs": 2000,
            "learning_rate": 0.001, "pose_lr": 0.0005, "distortion_lr": 0.0005,
        },
    }


def make_poses(n):
    """Forward-moving track: t_i = (0.02 i sin(0.3 i), 0, -0.15 i), yaw 0.01 i rad.  (n,4,4) c2w."""
    c2w = torch.eye(4).repeat(n, 1, 1)
    for i in range(n):
        a = 0.01 * i
        c, s = math.cos(a), math.sin(a)
        c2w[i, :3, :3] = torch.tensor([[c, 0.0, s], [0.0, 1.0, 0.0], [-s, 0.0, c]])
        c2w[i, :3, 3] = torch.tensor([0.02 * i * math.sin(0.3 * i), 0.0, -0.15 * i])
    return c2w


def make_scene(name="cfg1", frames=None, h=None, w=None, sparse_depth=False):
    """Returns dict: imgs (N,3,H,W), depths (N,H,W), depth_masks (N,H,W) bool, camera_mat (1,4,4),
    scale_mat (1,4,4), poses_gt (N,4,4)."""
    n0, h0, w0, _ = SHAPES[name]
    n, h, w = frames or n0, h or h0, w or w0
    fx = 725.0 * w / 1242.0  # V-KITTI focal scaled by image width (get_kittivirtual.py:12-14)
    fy = fx
    camera_mat = torch.diag(torch.tensor([2 * fx / w, -2 * fy / h, -1.0, 1.0])).unsqueeze(0)
    poses = make_poses(n)
    ys = torch.arange(h, dtype=torch.float32).view(1, 1, h, 1)
    xs = torch.arange(w, dtype=torch.float32).view(1, 1, 1, w)
    g1 = torch.Generator().manual_seed(1)
    g2 = torch.Generator().manual_seed(2)
    ph = poses[:, :3, 3].norm(dim=-1).view(n, 1, 1, 1)  # phase shift by the frame's pose
    ch = torch.arange(3, dtype=torch.float32).view(1, 3, 1, 1)
    imgs = 0.5 + 0.15 * torch.sin(2 * math.pi * (xs / w) * (1 + ch) + ph * 3 + ch) \
        + 0.15 * torch.cos(2 * math.pi * (ys / h) * (2 + ch) + ph) \
        + 0.15 * torch.sin(2 * math.pi * (xs / w + ys / h) * 1.5 + 0.7 * ch + 2 * ph)
    imgs = (imgs + 0.02 * torch.rand(n, 3, h, w, generator=g1)).clamp(0, 1).contiguous()
    depths = 1.5 + 0.8 * torch.sin(2 * math.pi * xs / w) * torch.cos(2 * math.pi * ys / h) + 0.3 * (ys / h)
    depths = (depths.expand(n, 1, h, w)[:, 0] + 0.01 * torch.rand(n, h, w, generator=g2)).contiguous()
    masks = torch.ones(n, h, w, dtype=torch.bool)
    if sparse_depth:  # [1,9,1,9] sparsity pattern (configs/V_KITTI/intersection_d5.yaml)
        keep = torch.zeros(h, w, dtype=torch.bool)
        keep[::10, ::10] = True
        masks = keep.unsqueeze(0).repeat(n, 1, 1)
        depths = depths * masks
    return {"imgs": imgs, "depths": depths, "depth_masks": masks, "camera_mat": camera_mat,
            "scale_mat": torch.eye(4).unsqueeze(0), "poses_gt": poses, "h": h, "w": w, "n": n}


def make_batch(scene, idx, ref_idx=None, dpt=True):
    """The batch dict the boundary consumes (dataloading/dataset.py:286-342)."""
    key = "img.dpt" if dpt else "img.depth"
    b = {
        "img": scene["imgs"][idx:idx + 1], "img.idx": torch.tensor([idx]),
        key: scene["depths"][idx:idx + 1], "img.depth_mask": scene["depth_masks"][idx:idx + 1],
        "img.camera_mat": scene["camera_mat"], "img.scale_mat": scene["scale_mat"],
        "img.pose_gt": scene["poses_gt"][idx], "img.gt_depths": scene["depths"][idx:idx + 1],
    }
    if ref_idx is not None:
        b["img.ref_imgs"] = scene["imgs"][ref_idx:ref_idx + 1]
        b["img.ref_dpts" if dpt else "img.ref_depths"] = scene["depths"][ref_idx:ref_idx + 1]
        b["img.ref_idxs"] = torch.tensor([ref_idx])
        b["img.ref_pose_gt"] = scene["poses_gt"][ref_idx]
    return b


def init_pose_params(n, seed=11, std=1e-2):
    g = torch.Generator().manual_seed(seed)
    return std * torch.randn(n, 3, generator=g), std * torch.randn(n, 3, generator=g)


def ray_choice(h, w, n_rays, seed=42):
    g = torch.Generator().manual_seed(seed)
    return torch.randperm(h * w, generator=g)[:n_rays]


def jitter(n_rays, n_samples=128, seed=43):
    g = torch.Generator().manual_seed(seed)
    return torch.rand(1, n_rays, n_samples, generator=g)


def init_mlp_state(seed=42, white_background=False):
    """Random-init weights with the reference's layer shapes, default nn.Linear init and the bias
    overrides of model/official_nerf.py:39-44.  Keys follow the reference state_dict."""
    import torch.nn as nn
    torch.manual_seed(seed)
    shapes = [("layers0.0", 63, 256), ("layers0.2", 256, 256), ("layers0.4", 256, 256), ("layers0.6", 256, 256),
              ("layers1.0", 319, 256), ("layers1.2", 256, 256), ("layers1.4", 256, 256), ("layers1.6", 256, 256),
              ("fc_density", 256, 1), ("fc_feature", 256, 256), ("rgb_layers.0", 283, 128), ("fc_rgb", 128, 3)]
    sd = {}
    for name, i, o in shapes:
        lin = nn.Linear(i, o)
        sd[name + ".weight"] = lin.weight.detach().clone()
        sd[name + ".bias"] = lin.bias.detach().clone()
    sd["fc_density.bias"] = torch.tensor([0.1])
    sd["fc_rgb.bias"] = torch.tensor([0.8] * 3 if white_background else [0.02] * 3)
    return sd
