"""Parity at BASELINE.json sizes on the benchmarked path.

cfg2: one full training step (20-frame 375 x 1242 scene, 4,096 rays x 128 samples, rgb + depth + Chamfer on 28,830-point clouds
+ reprojection, LearnPose, learned scale/shift) against tests/golden/cfg2_step.pt, which oracle/make_golden.py --cfg2 produced
by running the UNMODIFIED reference Trainer.compute_loss + backward on the same inputs (torch CPU fp32).  Checked: the fp32
CUDA-core path, and the bf16 tensor-core path through the CUDA-GRAPH step that bench.py times.
cfg5: nearest-neighbour indices at 100,000 x 100,000 against the reference's comp_closest_pts_idx_with_split on four
1,024-source slices (tests/golden/cfg5_chamfer.pt), bit-exact, clouds in raster and in random order, all three sweep modes.

Every test prints its measured worst-case errors (pytest -s); profiles/r2_parity_errors.txt is that output from a B200."""
import os
import sys

import pytest
import torch

from util import ROOT, build_models, dev, syn

pytestmark = pytest.mark.gpu

LOSS_KEYS = ("loss", "loss_rgb", "loss_depth", "l2_mean", "loss_pc", "loss_rgb_s")


def _setup():
    sys.path.insert(0, ROOT)
    from oracle.make_golden import cfg2_scene_setup
    return cfg2_scene_setup()


def _trainer(cfg, scene, r0, t0, scales, shifts, precision, flat_adam=False):
    import model as mdl
    tcfg = dict(cfg["training"], n_training_points=4096)
    nerf, renderer, net, pose, dist = build_models(cfg, scene["n"], scene, precision=precision, r0=r0, t0=t0, scales=scales, shifts=shifts)
    Adam = mdl.FlatAdam if flat_adam else torch.optim.Adam
    tr = mdl.Trainer(net, Adam(net.parameters(), lr=1e-3), tcfg, device=dev(), optimizer_pose=Adam(pose.parameters(), lr=5e-4),
                     pose_param_net=pose, optimizer_distortion=Adam(dist.parameters(), lr=5e-4), distortion_net=dist)
    return tr, nerf, pose, dist


def _rel(a, b):
    """scale-relative error: max |a - b| / max |b|."""
    a, b = a.detach().double().cpu().reshape(-1), b.detach().double().cpu().reshape(-1)
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def _rel_l2(a, b):
    a, b = a.detach().double().cpu().reshape(-1), b.detach().double().cpu().reshape(-1)
    return float((a - b).norm() / b.norm().clamp_min(1e-30))


def _stats(t):
    t = t.detach().double().cpu()
    return torch.stack([t.sum(), t.abs().sum(), (t * t).sum()])


def _compare(tag, rec, ld, nerf, pose, dist, loss_tol, grad_tol, mlp_tol, stat_tol):
    errs = {}
    for k in LOSS_KEYS:
        got, want = float(ld[k].reshape(-1)[0]), float(rec["loss_dict"][k].reshape(-1)[0])
        errs["loss/" + k] = abs(got - want) / max(abs(want), 1e-12)
        print("%s %-12s got %.8f  reference %.8f  rel %.2e" % (tag, k, got, want, errs["loss/" + k]))
        assert errs["loss/" + k] <= loss_tol, (tag, k, got, want)
    pairs = [("grad_r", pose.r.grad, rec["grad_r"]), ("grad_t", pose.t.grad, rec["grad_t"])]
    for gname, p, want in (("grad_scales", dist.global_scales, rec["grad_scales"]), ("grad_shifts", dist.global_shifts, rec["grad_shifts"])):
        if want is None:   # e.g. the pinned scale of the last frame: no gradient, so torch.optim.Adam skips the parameter
            assert p.grad is None, "%s: the reference leaves %s without gradient" % (tag, gname)
        else:
            pairs.append((gname, p.grad, want))
    for name, got, want in pairs:
        e = _rel(got, want)
        print("%s %-12s scale-relative err %.2e" % (tag, name, e))
        assert e <= grad_tol, (tag, name, e)
    P = dict(nerf.named_parameters())
    worst = 0.0
    for k, want in rec["mlp_grads"].items():
        e = _rel_l2(P[k].grad, want)
        worst = max(worst, e)
        print("%s d %-18s rel-L2 %.2e" % (tag, k, e))
        assert e <= mlp_tol, (tag, k, e)
    worst_stat = 0.0
    for k, want in rec["mlp_grad_stats"].items():
        got = _stats(P[k].grad)
        # sum |g| and sum g^2 of every one of the 24 gradient tensors (the plain sum cancels and is not a stable statistic)
        for j in (1, 2):
            e = abs(float(got[j]) - float(want[j])) / max(abs(float(want[j])), 1e-30)
            worst_stat = max(worst_stat, e)
            assert e <= stat_tol, (tag, k, j, float(got[j]), float(want[j]))
    print("%s worst MLP-gradient rel-L2 (7 tensors) %.2e, worst |g| / g^2 statistic over the 24 tensors %.2e" % (tag, worst, worst_stat))


@pytest.fixture(scope="module")
def cfg2_golden(golden):
    if not os.path.exists(os.path.join(ROOT, "tests", "golden", "cfg2_step.pt")):
        pytest.skip("tests/golden/cfg2_step.pt not generated (python oracle/make_golden.py --cfg2)")
    g = golden("cfg2_step")
    scene, r0, t0, scales, shifts, ray_idx, noise = _setup()
    assert int(ray_idx.sum()) == g["setup"]["ray_idx_checksum"] and abs(float(noise.double().sum()) - g["setup"]["noise_checksum"]) < 1e-6
    return g, scene, r0, t0, scales, shifts, ray_idx, noise


@pytest.mark.parametrize("case", ["mid", "last"])
def test_cfg2_step_fp32_vs_reference(cfg2_golden, case):
    g, scene, r0, t0, scales, shifts, ray_idx, noise = cfg2_golden
    rec = g["cases"][case]
    cfg = syn.default_cfg()
    tr, nerf, pose, dist = _trainer(cfg, scene, r0, t0, scales, shifts, "fp32")
    tr.fixed_ray_idx, tr.fixed_noise = ray_idx, noise
    batch = syn.make_batch(scene, rec["case"]["idx"], rec["case"]["ref"])
    ld = tr.compute_loss(batch, it=1, epoch=0, scheduling_start=cfg["training"]["scheduling_start"])
    ld["loss"].backward()
    torch.cuda.synchronize()
    # north_star: fp32 paths <= 1e-5 relative.  Measured on B200 (profiles/r2_parity_errors.txt): losses <= 8.3e-7, pose /
    # distortion gradients <= 1.6e-5 scale-relative, MLP gradients <= 2.3e-4 relative L2 except layers0.0.weight at 2.4e-3 (the
    # layer behind the 2^9 positional-encoding gain: an fp32 rounding of a sample position moves sin(512 x) by 512 ulp on BOTH
    # sides, which add up in a different order) -- asserted at 1e-5 / 1e-4 / 5e-3.
    _compare("cfg2/%s fp32" % case, rec, ld, nerf, pose, dist, loss_tol=1e-5, grad_tol=1e-4, mlp_tol=5e-3, stat_tol=1e-3)


@pytest.mark.parametrize("case", ["mid", "last"])
def test_cfg2_graphed_bf16_step_vs_reference(cfg2_golden, case):
    """The path bench.py times: bf16 tensor-core render inside the CUDA-graph step, FlatAdam.  Measured on B200 against the
    fp32 reference (profiles/r2_parity_errors.txt): losses <= 6.3e-6 relative, pose / distortion gradients <= 2.4e-5
    scale-relative, MLP gradients <= 2.9e-2 relative L2 (ReLU masks flip for pre-activations within bf16 rounding of zero; over
    4,096 rays the flips average out far below the 0.13 of the 64-ray fixture) -- asserted at 1e-4 / 1e-3 / 6e-2."""
    g, scene, r0, t0, scales, shifts, ray_idx, noise = cfg2_golden
    rec = g["cases"][case]
    cfg = syn.default_cfg()
    tr, nerf, pose, dist = _trainer(cfg, scene, r0, t0, scales, shifts, "bf16", flat_adam=True)
    tr.fixed_ray_idx, tr.fixed_noise = ray_idx.to(dev()), noise.to(dev())
    for o in (tr.optimizer, tr.optimizer_pose, tr.optimizer_distortion):
        o.param_groups[0]["lr"] = 0.0      # gradients and losses of THIS parameter state; the update itself is covered elsewhere
    batch = {k: (v.to(dev()) if torch.is_tensor(v) and k not in ("img.idx", "img.ref_idxs") else v)
             for k, v in syn.make_batch(scene, rec["case"]["idx"], rec["case"]["ref"]).items()}
    sched = cfg["training"]["scheduling_start"]
    ld = tr.train_step_graphed(batch, it=1, epoch=0, scheduling_start=sched)
    ld = tr.train_step_graphed(batch, it=2, epoch=0, scheduling_start=sched)    # a REPLAY of the captured graph
    torch.cuda.synchronize()
    _compare("cfg2/%s bf16 graphed" % case, rec, ld, nerf, pose, dist, loss_tol=1e-4, grad_tol=1e-3, mlp_tol=6e-2, stat_tol=3e-2)


@pytest.mark.parametrize("case", ["mid", "last"])
def test_cfg2_chamfer_indices_and_render_vs_reference(cfg2_golden, case):
    """The 2 x 28,830 nearest-neighbour indices of the step's aligned clouds, as the reference's own function returned them, and
    the rendered rgb / depth of the 4,096 rays on both precision paths."""
    from model import _ops
    g, scene, r0, t0, scales, shifts, ray_idx, noise = cfg2_golden
    rec = g["cases"][case]
    X, Y = rec["X"].to(dev()), rec["Y"].to(dev())
    ixy, iyx = _ops.chamfer_nn(X, Y).cpu(), _ops.chamfer_nn(Y, X).cpu()
    for name, got, want in (("X->Y", ixy, rec["nn_idx_xy"].long()), ("Y->X", iyx, rec["nn_idx_yx"].long())):
        diff = got != want
        if diff.any():
            # only documented fp32 distance ties may differ (the reference's argmin over a materialised matrix and ours return
            # the first minimum; a differing index must be an exact tie)
            a, b = (X, Y) if name == "X->Y" else (Y, X)
            i = torch.nonzero(diff).reshape(-1)
            d_got = torch.linalg.norm(a.cpu()[i] - b.cpu()[got[i]], dim=1)
            d_want = torch.linalg.norm(a.cpu()[i] - b.cpu()[want[i]], dim=1)
            assert torch.equal(d_got, d_want) and bool((got[i] < want[i]).all()), (name, int(diff.sum()))
        print("cfg2/%s Chamfer %s: %d of %d indices differ from the reference's" % (case, name, int(diff.sum()), got.numel()))
        assert int(diff.sum()) == 0
    cfg = syn.default_cfg()
    # measured on B200: fp32 rgb 1.8e-7 abs / depth 3.2e-7 rel; bf16 rgb 1.2e-4 abs / depth 3.4e-5 rel (north_star: <= 1e-2 abs rgb)
    for precision, tol in (("fp32", 1e-5), ("bf16", 1e-3)):
        tr, nerf, pose, dist = _trainer(cfg, scene, r0, t0, scales, shifts, precision)
        tr.fixed_ray_idx, tr.fixed_noise = ray_idx, noise
        seen = {}
        fwd = tr.model.renderer.forward

        def rec_fwd(*a, _f=fwd, **k):
            out = _f(*a, **k)
            seen.update(out)
            return out

        tr.model.renderer.forward = rec_fwd
        with torch.no_grad():
            tr.compute_loss(syn.make_batch(scene, rec["case"]["idx"], rec["case"]["ref"]), it=1, epoch=0,
                            scheduling_start=cfg["training"]["scheduling_start"])
        e_rgb = float((seen["rgb"].cpu() - rec["rgb"]).abs().max())
        e_d = _rel(seen["depth_pred_dense"][seen["mask"]], rec["depth_pred"])
        e_g = _rel(seen["depth_gt_dense"][seen["mask"]], rec["depth_gt"])
        print("cfg2/%s %s render: rgb max abs %.2e, depth_pred rel %.2e, depth_gt rel %.2e" % (case, precision, e_rgb, e_d, e_g))
        assert e_rgb <= tol and e_d <= tol and e_g <= 1e-5


@pytest.fixture(params=["tcgen05", "cuda_cores", "mixed"])
def chamfer_pass1(request):
    from model import _cabi as K
    old = K.lib().nnb_debug_chamfer_pass1({"tcgen05": 0, "cuda_cores": 1, "mixed": 2}[request.param])
    yield request.param
    K.lib().nnb_debug_chamfer_pass1(old)


@pytest.mark.parametrize("order", ["raster", "random"])
def test_cfg5_chamfer_100k_indices_bit_exact(golden, order, chamfer_pass1):
    """BASELINE.json cfg5: 100,000 x 100,000.  The whole problem runs on the GPU; the four 1,024-source slices the reference
    evaluated (model/losses.py:129-144 on pts_src[:, a:b]) must agree index for index."""
    from model import _ops
    sys.path.insert(0, ROOT)
    from oracle.make_golden import cfg5_clouds
    g = golden("cfg5_chamfer")
    a, b = cfg5_clouds()[order]
    idx, dist = _ops.chamfer_nn(a.to(dev()), b.to(dev()), want_dist=True)
    idx = idx.cpu()
    assert int(idx.min()) >= 0 and int(idx.max()) < 100000
    n_bad = 0
    for s0, want in g[order].items():
        n_bad += int((idx[s0:s0 + g["slice_len"]] != want.long()).sum())
    print("cfg5 %s %s: %d of %d checked indices differ" % (order, chamfer_pass1, n_bad, 4 * g["slice_len"]))
    assert n_bad == 0
    # size-independent property on ALL 100,000 sources: the returned distance is the distance to the returned index, and no
    # point of a random probe set of destinations is closer
    d_chk = torch.linalg.norm(a - b[idx], dim=1)
    torch.testing.assert_close(dist.cpu(), d_chk, rtol=1e-6, atol=1e-7)
    probe = torch.randint(0, 100000, (512,), generator=torch.Generator().manual_seed(1))
    d_probe = torch.cdist(a.to(dev()), b[probe].to(dev()), compute_mode="donot_use_mm_for_euclid_dist").min(dim=1).values.cpu()
    assert bool((d_chk <= d_probe * (1 + 1e-6) + 1e-7).all())
