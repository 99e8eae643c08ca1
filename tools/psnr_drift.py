"""PSNR drift between the fp32 CUDA-core path and the bf16 tensor-core path over a training run (north_star: "<= 0.05 dB PSNR
drift over 1k steps if bf16 tensor cores are used").  Same scene, same initial weights, same ray choice and jitter per step
(seeded), same Adam; PSNR of a held-out full-frame render every `every` steps.
usage: python tools/psnr_drift.py [steps] [rays] [every]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "my-nope-nerf_b200")):
    sys.path.insert(0, p)
import torch
import synthetic_scene as syn
import model as mdl
from model import _ops

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
R = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
every = int(sys.argv[3]) if len(sys.argv) > 3 else 100
dev = torch.device("cuda:0")


def run(precision):
    cfg = syn.default_cfg()
    scene = syn.make_scene("tiny")
    n, h, w = scene["n"], scene["h"], scene["w"]
    nerf = mdl.OfficialStaticNerf(cfg)
    nerf.load_state_dict(syn.init_mlp_state(42))
    net = mdl.get_model(mdl.Renderer(nerf, dict(cfg["rendering"], precision=precision), device=dev), cfg, device=dev).to(dev)
    pose = mdl.LearnPose(n, True, True, cfg, init_c2w=scene["poses_gt"].clone()).to(dev)
    dist = mdl.Learn_Distortion(n, True, True, cfg).to(dev)
    tcfg = dict(cfg["training"], n_training_points=R)
    tr = mdl.Trainer(net, mdl.FlatAdam(net.parameters(), lr=tcfg["learning_rate"]), tcfg, device=dev,
                     optimizer_pose=mdl.FlatAdam(pose.parameters(), lr=tcfg["pose_lr"]), pose_param_net=pose,
                     optimizer_distortion=mdl.FlatAdam(dist.parameters(), lr=tcfg["distortion_lr"]), distortion_net=dist)
    rs = mdl.ResidentScene([syn.make_batch(scene, i) for i in range(n)], dev)
    ex = mdl.Extract_Images(net.renderer, {"extract_images": {"resolution": [h, w]}}, device=dev, render_type=tr.rendering_technique)
    cam, scl = scene["camera_mat"].to(dev).reshape(1, 4, 4), torch.eye(4, device=dev).reshape(1, 4, 4)
    psnr = []
    sched = cfg["training"]["scheduling_start"]
    for it in range(steps + 1):
        if it % every == 0:
            with torch.no_grad():     # mean PSNR over every frame of the sequence
                net.eval()
                vals = []
                for f in range(n):
                    wm = _ops.Mat4Inverse.apply(pose(f)).unsqueeze(0)
                    rgb, _ = ex.render_frame(cam, wm, scl, h, w)
                    mse = float(((rgb.permute(2, 0, 1) - scene["imgs"][f].to(dev)) ** 2).mean())
                    vals.append(-10.0 * torch.log10(torch.tensor(max(mse, 1e-10))).item())
                net.train()
                psnr.append(sum(vals) / n)
        if it == steps:
            break
        torch.manual_seed(1000 + it)          # same ray choice and jitter on both paths
        f = it % n
        tr.train_step(rs.batch(f, f + 1 if f < n - 1 else f - 1), it=it, epoch=0, scheduling_start=sched)
    return psnr


if __name__ == "__main__":
    a, a2, b = run("fp32"), run("fp32"), run("bf16")
    # the fp32 path accumulates weight gradients with atomics (split-K): two fp32 runs already diverge chaotically -- that is the
    # noise floor the bf16 drift has to be read against
    print("step  PSNR fp32   fp32 again  PSNR bf16   fp32-fp32 dB   bf16-fp32 dB")
    for i, (x, x2, y) in enumerate(zip(a, a2, b)):
        print("%5d  %9.4f  %9.4f  %9.4f  %+12.4f  %+12.4f" % (i * every, x, x2, y, x2 - x, y - x))
    k = max(1, len(a) // 3)
    m = lambda v: sum(v[-k:]) / k
    print("mean PSNR over the last %d evaluations: fp32 %.3f, fp32 again %.3f, bf16 %.3f dB -> run-to-run %.3f dB, bf16 drift %.3f dB"
          % (k, m(a), m(a2), m(b), m(a2) - m(a), m(b) - m(a)))
