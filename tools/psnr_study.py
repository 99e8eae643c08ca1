"""PSNR drift of the bf16 tensor-core path against the fp32 CUDA-core path, as a PAIRED multi-seed study (north_star: "<= 0.05 dB
PSNR drift over 1k steps if bf16 tensor cores are used").

BASELINE.json cfg1 scene (8 frames, 120 x 160), 1,024 rays x 128 samples per step, full loss set, three Adam optimisers,
`steps` steps.  A seed fixes the MLP initialisation and, step by step, the ray choice and the jitter, identically on both arms
(eager step: every step starts with torch.manual_seed).  Per seed: final PSNR = mean over the last three evaluations of the mean
PSNR of full-frame renders of all 8 frames.  Reported: per-arm mean +- std, the paired difference bf16 - fp32 with its standard
error and 95 % interval (t distribution), and -- the noise floor -- the same statistic for fp32 against a REPEAT of fp32 with the
same seed (the fp32 path's split-K weight gradient adds with atomics, so even identical seeds separate).

usage: python tools/psnr_study.py [steps=1000] [seeds=5] [every=100] [first_seed=42]  -> markdown on stdout"""
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "my-nope-nerf_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402
import synthetic_scene as syn  # noqa: E402
import model as mdl  # noqa: E402
from model import _ops  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
n_seeds = int(sys.argv[2]) if len(sys.argv) > 2 else 5
every = int(sys.argv[3]) if len(sys.argv) > 3 else 100
seed0 = int(sys.argv[4]) if len(sys.argv) > 4 else 42
R = 1024
dev = torch.device("cuda:0")
T95 = {2: 12.706, 3: 4.303, 4: 3.182, 5: 2.776, 6: 2.571, 7: 2.447, 8: 2.365, 9: 2.306, 10: 2.262}


def run(precision, seed):
    cfg = syn.default_cfg()
    scene = syn.make_scene("cfg1")
    n, h, w = scene["n"], scene["h"], scene["w"]
    nerf = mdl.OfficialStaticNerf(cfg)
    nerf.load_state_dict(syn.init_mlp_state(seed))
    net = mdl.get_model(mdl.Renderer(nerf, dict(cfg["rendering"], precision=precision), device=dev), cfg, device=dev).to(dev)
    pose = mdl.LearnPose(n, True, True, cfg, init_c2w=scene["poses_gt"].clone()).to(dev)
    dist = mdl.Learn_Distortion(n, True, True, cfg).to(dev)
    tcfg = dict(cfg["training"], n_training_points=R)
    tr = mdl.Trainer(net, mdl.FlatAdam(net.parameters(), lr=tcfg["learning_rate"]), tcfg, device=dev,
                     optimizer_pose=mdl.FlatAdam(pose.parameters(), lr=tcfg["pose_lr"]), pose_param_net=pose,
                     optimizer_distortion=mdl.FlatAdam(dist.parameters(), lr=tcfg["distortion_lr"]), distortion_net=dist)
    tr.use_graph = False          # eager: torch.manual_seed before a step fixes its draws identically on both arms
    rs = mdl.ResidentScene([syn.make_batch(scene, i) for i in range(n)], dev)
    ex = mdl.Extract_Images(net.renderer, {"extract_images": {"resolution": [h, w]}}, device=dev, render_type=tr.rendering_technique)
    cam, scl = scene["camera_mat"].to(dev).reshape(1, 4, 4), torch.eye(4, device=dev).reshape(1, 4, 4)
    imgs = scene["imgs"].to(dev)
    psnr = []
    sched = cfg["training"]["scheduling_start"]
    for it in range(steps + 1):
        if it % every == 0:
            with torch.no_grad():
                net.eval()
                vals = []
                for f in range(n):
                    wm = _ops.Mat4Inverse.apply(pose(f)).unsqueeze(0)
                    rgb, _ = ex.render_frame(cam, wm, scl, h, w)
                    mse = float(((rgb.permute(2, 0, 1) - imgs[f]) ** 2).mean())
                    vals.append(-10.0 * math.log10(max(mse, 1e-10)))
                net.train()
                psnr.append(sum(vals) / n)
        if it == steps:
            break
        torch.manual_seed(seed * 1000003 + it)
        f = it % n
        tr.train_step(rs.batch(f, f + 1 if f < n - 1 else f - 1), it=it, epoch=0, scheduling_start=sched)
    return psnr


def final(p):
    return sum(p[-3:]) / 3.0


def stats(v):
    n = len(v)
    m = sum(v) / n
    sd = math.sqrt(sum((x - m) ** 2 for x in v) / (n - 1)) if n > 1 else 0.0
    return m, sd, sd / math.sqrt(n), T95.get(n, 2.0) * sd / math.sqrt(n)


if __name__ == "__main__":
    seeds = list(range(seed0, seed0 + n_seeds))
    rows = []
    for s in seeds:
        a, a2, b = run("fp32", s), run("fp32", s), run("bf16", s)
        rows.append((s, a, a2, b))
        sys.stderr.write("seed %d done: fp32 %.3f, fp32 repeat %.3f, bf16 %.3f\n" % (s, final(a), final(a2), final(b)))
    print("# PSNR drift study: bf16 tensor-core path vs fp32 CUDA-core path (tools/psnr_study.py %d %d %d, B200)\n" % (steps, n_seeds, every))
    print("cfg1 scene (8 frames 120 x 160), 1,024 rays x 128 samples per step, full loss set, 3 x FlatAdam, %d steps; PSNR = mean over the 8 "
          "frames of full-frame renders; final = mean of the last three evaluations (steps %d, %d, %d).\n" % (steps, steps - 2 * every, steps - every, steps))
    print("| seed | start | fp32 final | fp32 repeat final | bf16 final | fp32 repeat - fp32 | bf16 - fp32 |")
    print("|---|---|---|---|---|---|---|")
    for s, a, a2, b in rows:
        print("| %d | %.3f | %.3f | %.3f | %.3f | %+.3f | %+.3f |" % (s, a[0], final(a), final(a2), final(b), final(a2) - final(a), final(b) - final(a)))
    fa, fa2, fb = [final(r[1]) for r in rows], [final(r[2]) for r in rows], [final(r[3]) for r in rows]
    print()
    for name, v in (("fp32", fa), ("fp32 repeat", fa2), ("bf16", fb)):
        m, sd, se, ci = stats(v)
        print("* %s final PSNR: mean %.3f dB, std over seeds %.3f dB" % (name, m, sd))
    for name, d in (("fp32 repeat - fp32 (same seed: the fp32 path's own run-to-run noise)", [x - y for x, y in zip(fa2, fa)]),
                    ("bf16 - fp32 (paired by seed)", [x - y for x, y in zip(fb, fa)])):
        m, sd, se, ci = stats(d)
        print("* %s: mean %+.3f dB, std %.3f dB, standard error %.3f dB, 95 %% interval [%+.3f, %+.3f] dB, largest |difference| %.3f dB"
              % (name, m, sd, se, m - ci, m + ci, max(abs(x) for x in d)))
    print("\nPSNR at every evaluation (seed: fp32 / fp32 repeat / bf16):\n")
    print("| step | " + " | ".join("seed %d" % r[0] for r in rows) + " |")
    print("|---|" + "---|" * len(rows))
    for i in range(len(rows[0][1])):
        print("| %d | " % (i * every) + " | ".join("%.2f / %.2f / %.2f" % (r[1][i], r[2][i], r[3][i]) for r in rows) + " |")
