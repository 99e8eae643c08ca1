"""Generate tests/golden/*.pt by running the UNMODIFIED reference (imported read-only from
/root/reference through oracle/ref_harness.py) on seeded synthetic inputs.

Run in the build container only:   python oracle/make_golden.py
The fixtures travel to the GPU box; the reference does not.  TEST INFRASTRUCTURE ONLY.
"""
import contextlib
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from oracle import ref_harness  # noqa: E402
import synthetic_scene as syn  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


@contextlib.contextmanager
def injected_draws(noise=None, perm=None):
    """Replace torch.rand / torch.randperm for the duration of one reference call so that the
    reference consumes OUR jitter (model/rendering.py:190) and OUR ray choice (model/training.py:277)."""
    o_rand, o_perm = torch.rand, torch.randperm

    def rand(*shape, **kw):
        assert noise is not None and tuple(shape) == tuple(noise.shape), (shape, None if noise is None else noise.shape)
        return noise.clone()

    def randperm(n, **kw):
        assert perm is not None
        return perm.clone()

    torch.rand, torch.randperm = rand, randperm
    try:
        yield
    finally:
        torch.rand, torch.randperm = o_rand, o_perm


def build_ref_model(ref, cfg, sd):
    nerf = ref.model.OfficialStaticNerf(cfg)
    nerf.load_state_dict(sd)
    renderer = ref.model.Renderer(nerf, cfg["rendering"], device=torch.device("cpu"))
    net = ref.network.nope_nerf(cfg, renderer, None, torch.device("cpu"))
    return nerf, renderer, net


def stage_vectors(ref, cfg):
    g = torch.Generator().manual_seed(5)
    out = {"torch_version": torch.__version__}
    # --- SE(3) exp / pose / distortion (model/common.py:277-310, model/poses.py, model/distortions.py)
    rs = torch.cat([torch.zeros(1, 3), 1e-2 * torch.randn(3, 3, generator=g), torch.randn(2, 3, generator=g),
                    torch.tensor([[3.0, 0.1, -0.2]])])
    ts = torch.randn(rs.shape[0], 3, generator=g)
    init = syn.make_poses(rs.shape[0])
    c2w, c2w_init, g_r, g_t = [], [], [], []
    probe = torch.randn(4, 4, generator=g)
    for r, t, ini in zip(rs, ts, init):
        r = r.clone().requires_grad_(True)
        t = t.clone().requires_grad_(True)
        m = ref.common.make_c2w(r, t)
        mi = m @ ini
        (mi * probe).sum().backward()
        c2w.append(m.detach())
        c2w_init.append(mi.detach())
        g_r.append(r.grad.clone())
        g_t.append(t.grad.clone())
    out["pose"] = {"r": rs, "t": ts, "init": init, "probe": probe, "c2w": torch.stack(c2w),
                   "c2w_init": torch.stack(c2w_init), "grad_r": torch.stack(g_r), "grad_t": torch.stack(g_t)}
    dn = ref.model.Learn_Distortion(4, True, True, cfg)
    with torch.no_grad():
        dn.global_scales.copy_(torch.tensor([[1.3], [0.005], [0.7], [2.0]]))
        dn.global_shifts.copy_(torch.tensor([[0.1], [-0.2], [0.3], [0.4]]))
    out["distortion"] = {"scales": dn.global_scales.detach().clone(), "shifts": dn.global_shifts.detach().clone(),
                         "out": [tuple(float(v) for v in dn(i)) for i in range(4)]}
    # --- pixel grid (model/common.py:13-40)
    loc, scaled = ref.common.arange_pixels((5, 7))
    out["pixel_grid"] = {"h": 5, "w": 7, "loc": loc, "scaled": scaled}
    # --- positional encoding (model/official_nerf.py:99-119): far-plane-sized arguments included
    x = torch.cat([torch.randn(64, 3, generator=g), 10.0 * (torch.rand(64, 3, generator=g) - 0.5) * 2])
    out["posenc"] = {"x": x, "enc10": ref.official_nerf.encode_position(x, 10, True),
                     "enc4": ref.official_nerf.encode_position(x, 4, True)}
    # --- compositing semantics probes (model/rendering.py:116-126)
    a = torch.rand(16, 128, generator=g) * 0.2
    a[0, 5] = 1.0
    a[1] = 0.0
    c = torch.rand(16, 128, 3, generator=g)
    z = torch.sort(torch.rand(16, 128, generator=g) * 10)[0]
    a_ = a.clone().requires_grad_(True)
    c_ = c.clone().requires_grad_(True)
    w = a_ * torch.cumprod(torch.cat([torch.ones(16, 1), 1. - a_ + 1e-6], -1), -1)[:, :-1]
    rgb = torch.sum(w.unsqueeze(-1) * c_, dim=-2)
    dist = torch.sum(w * z, dim=-1)
    pr, pd = torch.randn(16, 3, generator=g), torch.randn(16, generator=g)
    ((rgb * pr).sum() + (dist * pd).sum()).backward()
    out["composite"] = {"alpha": a, "rgb_s": c, "z": z, "weights": w.detach(), "rgb": rgb.detach(), "dist": dist.detach(),
                        "probe_rgb": pr, "probe_dist": pd, "grad_alpha": a_.grad.clone(), "grad_rgb_s": c_.grad.clone()}
    return out


def loss_vectors(ref, cfg):
    g = torch.Generator().manual_seed(6)
    L = ref.losses.Loss(cfg["training"])
    out = {}
    pred = torch.rand(1, 50, 3, generator=g)
    gt = torch.rand(1, 50, 3, generator=g)
    out["rgb"] = {"pred": pred, "gt": gt, "l1": L.get_rgb_full_loss(pred, gt, "l1"), "l2": L.get_rgb_full_loss(pred, gt, "l2"),
                  "mse": torch.nn.functional.mse_loss(pred, gt)}
    for n in (50, 51):  # even (lower median) and odd counts
        dp = (torch.rand(n, generator=g) * 5 + 0.5).requires_grad_(True)
        dg = (torch.rand(n, generator=g) * 5 + 0.5).requires_grad_(True)
        L.depth_loss_type = "l1"
        l1 = L.get_depth_loss(dp, dg)
        L.depth_loss_type = "invariant"
        inv = L.get_depth_loss(dp, dg)
        inv.backward()
        out["depth_%d" % n] = {"pred": dp.detach().clone(), "gt": dg.detach().clone(), "l1": l1.detach(), "inv": inv.detach(),
                               "inv_grad_pred": dp.grad.clone(), "inv_grad_gt": dg.grad.clone()}
    # Chamfer: random clouds, a lattice (exact ties -> first index), a coincident point (zero subgradient)
    X = torch.rand(3, 300, generator=g) * 2 - 1
    Y = torch.rand(3, 257, generator=g) * 2 - 1
    Y[:, 7] = X[:, 11]
    lat = torch.stack(torch.meshgrid(torch.arange(4.), torch.arange(4.), torch.arange(4.), indexing="ij")).reshape(3, -1)
    latq = lat[:, ::3] + 0.5
    Xr = X.clone().requires_grad_(True)
    Yr = Y.clone().requires_grad_(True)
    e = L.get_pc_loss(Xr.t().unsqueeze(0), Yr.t().unsqueeze(0))
    e.backward()
    out["chamfer"] = {"X": X, "Y": Y, "idx_xy": L.comp_closest_pts_idx_with_split(X, Y), "idx_yx": L.comp_closest_pts_idx_with_split(Y, X),
                      "loss": e.detach(), "grad_X": Xr.grad.clone(), "grad_Y": Yr.grad.clone(),
                      "lat_src": latq, "lat_dst": lat, "lat_idx": L.comp_closest_pts_idx_with_split(latq, lat)}
    r1 = torch.rand(1, 6, 8, 3, generator=g)
    r2 = torch.rand(1, 6, 8, 3, generator=g) * 2.5
    vm = torch.rand(1, 6, 8, 1, generator=g) > 0.3
    out["rgb_s"] = {"rgb1": r1, "rgb2": r2, "valid": vm, "loss": L.get_rgb_s_loss(r1, r2, vm),
                    "loss_empty": torch.tensor(0.0)}
    return out


def flat_stats(t):
    return torch.stack([t.sum(), t.abs().sum(), (t * t).sum()])


def render_and_step_vectors(ref, cfg):
    out = {}
    scene = syn.make_scene("tiny")
    n, h, w, R = scene["n"], scene["h"], scene["w"], syn.SHAPES["tiny"][3]
    sd = syn.init_mlp_state(42)
    nerf, renderer, net = build_ref_model(ref, cfg, sd)
    out["mlp_checksum"] = {k: flat_stats(v) for k, v in sd.items()}
    ray_idx = syn.ray_choice(h, w, R)
    noise = syn.jitter(R, 128)
    r0, t0 = syn.init_pose_params(n)
    # --- MLP forward alone (model/official_nerf.py:69-96)
    g = torch.Generator().manual_seed(9)
    pts = (torch.rand(256, 3, generator=g) - 0.5) * 8
    vd = torch.nn.functional.normalize(torch.randn(256, 3, generator=g), dim=-1)
    with torch.no_grad():
        rgb, alpha = nerf(pts, vd, return_addocc=True)
    out["mlp"] = {"pts": pts, "viewdir": vd, "rgb": rgb, "alpha": alpha}
    # --- Renderer.forward train (jitter injected) and eval (no jitter) (model/rendering.py:22-168)
    pose_net = ref.model.LearnPose(n, True, True, cfg, init_c2w=scene["poses_gt"].clone())
    with torch.no_grad():
        pose_net.r.copy_(r0)
        pose_net.t.copy_(t0)
    c2w = pose_net(1)
    world_mat = torch.inverse(c2w).unsqueeze(0).detach()
    _, p_full = ref.common.arange_pixels((h, w))
    p = p_full[:, ray_idx]
    depth_sel = scene["depths"][1].reshape(1, -1, 1)[:, ray_idx].clone()
    depth_sel[0, 3, 0] = 0.0  # one invalid-depth ray exercises the mask path
    with injected_draws(noise=noise):
        o_tr = renderer(p, depth_sel, scene["camera_mat"], world_mat, scene["scale_mat"], "nope_nerf", add_noise=True, eval_=False)
    with torch.no_grad():
        o_ev = renderer(p, depth_sel, scene["camera_mat"], world_mat, scene["scale_mat"], "nope_nerf", add_noise=False, eval_=True)
    out["render"] = {"ray_idx": ray_idx, "noise": noise, "pixels": p, "depth": depth_sel, "world_mat": world_mat,
                     "train": {k: (v.detach() if torch.is_tensor(v) else v) for k, v in o_tr.items()},
                     "eval": {k: (v.detach() if torch.is_tensor(v) else v) for k, v in o_ev.items()}}
    # --- full Trainer.compute_loss + backward (model/training.py:214-416), several branches
    cases = {
        "l1_mid": dict(idx=1, ref=2, depth_loss_type="l1", epoch=0),
        "inv_mid": dict(idx=1, ref=2, depth_loss_type="invariant", epoch=0),
        "l1_last": dict(idx=n - 1, ref=n - 2, depth_loss_type="l1", epoch=0),
        "l2_rgb_only": dict(idx=0, ref=1, depth_loss_type="l1", epoch=10 ** 6),
    }
    if os.environ.get("NNB_GOLDEN_EXTRA_CASES"):   # appended later without regenerating the cases above
        cases = {"shift_first": dict(idx=2, ref=3, depth_loss_type="l1", epoch=0, shift_first=True)}
    out["step"] = {}
    for name, c in cases.items():
        tcfg = dict(cfg["training"])
        tcfg["depth_loss_type"] = c["depth_loss_type"]
        tcfg["n_training_points"] = R
        tcfg["shift_first"] = c.get("shift_first", False)
        nerf, renderer, net = build_ref_model(ref, cfg, sd)
        pose_net = ref.model.LearnPose(n, True, True, cfg, init_c2w=scene["poses_gt"].clone())
        dist_net = ref.model.Learn_Distortion(n, True, True, cfg)
        with torch.no_grad():
            pose_net.r.copy_(r0)
            pose_net.t.copy_(t0)
            dist_net.global_scales.copy_(1.0 + 0.05 * torch.arange(n, dtype=torch.float32).view(n, 1))
            dist_net.global_shifts.copy_(0.01 * torch.arange(n, dtype=torch.float32).view(n, 1))
        trainer = ref.model.Trainer(net, None, tcfg, device=torch.device("cpu"), pose_param_net=pose_net,
                                    distortion_net=dist_net)
        batch = syn.make_batch(scene, c["idx"], c["ref"])
        with injected_draws(noise=noise, perm=torch.cat([ray_idx, torch.zeros(h * w - R, dtype=torch.long)])):
            ld = trainer.compute_loss(batch, it=1, epoch=c["epoch"], scheduling_start=tcfg["scheduling_start"],
                                      out_render_path="/tmp")
        ld["loss"].backward()
        rec = {"case": c, "loss_dict": {k: v.detach().clone().reshape(-1) for k, v in ld.items()},
               "grad_r": pose_net.r.grad.clone(), "grad_t": pose_net.t.grad.clone(),
               "grad_scales": dist_net.global_scales.grad.clone() if dist_net.global_scales.grad is not None else None,
               "grad_shifts": dist_net.global_shifts.grad.clone() if dist_net.global_shifts.grad is not None else None,
               "mlp_grad_stats": {k: flat_stats(v.grad) for k, v in nerf.named_parameters()}}
        if name == "l1_mid":
            rec["mlp_grads"] = {k: v.grad.clone() for k, v in nerf.named_parameters()}
        out["step"][name] = rec
        print(name, {k: float(v.reshape(-1)[0]) for k, v in ld.items()})
    out["step_setup"] = {"scales": 1.0 + 0.05 * torch.arange(n, dtype=torch.float32).view(n, 1),
                         "shifts": 0.01 * torch.arange(n, dtype=torch.float32).view(n, 1), "r": r0, "t": t0}
    return out


def variant_vectors(ref, cfg0):
    """Renderer.forward on the non-default branches: the LLFF configuration (configs/LLFF/fern.yaml:
    sample_option ndc + dist_alpha + depth_range [0,1]), dist_alpha with uniform sampling, and white background + relu
    density + use_ray_dir False.  Forward outputs, and gradients of a probed scalar w.r.t. the MLP, world_mat, camera_mat."""
    import copy
    out = {"torch_version": torch.__version__}
    scene = syn.make_scene("tiny")
    n, h, w = scene["n"], scene["h"], scene["w"]
    R = 48
    sd = syn.init_mlp_state(42)
    g = torch.Generator().manual_seed(21)
    ray_idx = torch.randperm(h * w, generator=g)[:R]
    noise = torch.rand(1, R, 128, generator=g)
    pr, pd = torch.randn(1, R, 3, generator=g) / R, torch.randn(R, generator=g) / R
    r0, t0 = syn.init_pose_params(n)
    pose_net = ref.model.LearnPose(n, True, True, cfg0, init_c2w=scene["poses_gt"].clone())
    with torch.no_grad():
        pose_net.r.copy_(r0)
        pose_net.t.copy_(t0)
    world_mat0 = torch.inverse(pose_net(1)).unsqueeze(0).detach()
    _, p_full = ref.common.arange_pixels((h, w))
    p = p_full[:, ray_idx]
    depth_sel = scene["depths"][1].reshape(1, -1, 1)[:, ray_idx].clone()
    variants = {
        "llff": dict(rendering=dict(sample_option="ndc", dist_alpha=True, depth_range=[0.0, 1.0])),
        "dist_alpha_uniform": dict(rendering=dict(dist_alpha=True)),
        "white_relu_nodir": dict(rendering=dict(white_background=True, use_ray_dir=False), model=dict(occ_activation="relu")),
    }
    out["inputs"] = {"ray_idx": ray_idx, "noise": noise, "pixels": p, "depth": depth_sel, "world_mat": world_mat0,
                     "camera_mat": scene["camera_mat"], "probe_rgb": pr, "probe_dist": pd}
    for name, over in variants.items():
        cfg = copy.deepcopy(cfg0)
        for sec, kv in over.items():
            cfg[sec].update(kv)
        nerf, renderer, net = build_ref_model(ref, cfg, sd)
        wm = world_mat0.clone().requires_grad_(True)
        cm = scene["camera_mat"].clone().requires_grad_(True)
        with injected_draws(noise=noise):
            o = renderer(p, depth_sel, cm, wm, scene["scale_mat"], "nope_nerf", add_noise=True, eval_=False)
        ((o["rgb"] * pr).sum() + (o["depth_pred"] * pd[:o["depth_pred"].shape[0]]).sum()).backward()
        with torch.no_grad():
            o_ev = renderer(p, depth_sel, scene["camera_mat"], world_mat0, scene["scale_mat"], "nope_nerf", add_noise=False, eval_=True)
        out[name] = {"override": over,
                     "train": {k: (v.detach().clone() if torch.is_tensor(v) else v) for k, v in o.items()},
                     "eval": {k: (v.detach().clone() if torch.is_tensor(v) else v) for k, v in o_ev.items()},
                     "grad_world_mat": wm.grad.clone(), "grad_camera_mat": cm.grad.clone(),
                     "mlp_grad_stats": {k: flat_stats(v.grad) for k, v in nerf.named_parameters()},
                     "mlp_grads": {k: v.grad.clone() for k, v in nerf.named_parameters()
                                   if k in ("layers0.0.weight", "fc_density.weight", "fc_density.bias", "fc_rgb.weight", "rgb_layers.0.bias")}}
        print(name, float(o["rgb"].sum()), float(o["depth_pred"].sum()))
    return out


def geometry_state(sd):
    """Random-init weights put the whole field on one side of the occupancy threshold (raw density 0.096 +- 0.004): scale the
    density head so that the field crosses tau = 0.5 inside the sphere (raw = 0 <=> occupancy 0.5)."""
    sd = {k: v.clone() for k, v in sd.items()}
    sd["fc_density.weight"] = sd["fc_density.weight"] * 200.0
    sd["fc_density.bias"] = torch.tensor([0.86])
    return sd


def geometry_vectors(ref, cfg):
    """OfficialStaticNerf.gradient, Renderer.ray_marching (+ secant) and Renderer.phong_renderer of the
    reference on a field with surfaces (geometry_state)."""
    out = {"torch_version": torch.__version__}
    scene = syn.make_scene("tiny")
    h, w = scene["h"], scene["w"]
    sd = geometry_state(syn.init_mlp_state(42))
    nerf, renderer, net = build_ref_model(ref, cfg, sd)
    g = torch.Generator().manual_seed(31)
    pts = (torch.rand(300, 3, generator=g) - 0.5) * 6
    grad = nerf.gradient(pts.clone(), 0).detach()
    with torch.no_grad():
        _, raw = nerf.infer_occ(pts)
        occ = nerf(pts, only_occupancy=True)
    out["gradient"] = {"pts": pts, "grad": grad, "raw": raw, "occ": occ}
    R = 96
    ray_idx = torch.randperm(h * w, generator=g)[:R]
    _, p_full = ref.common.arange_pixels((h, w))
    p = p_full[:, ray_idx]
    world_mat = torch.inverse(scene["poses_gt"][1]).unsqueeze(0)
    cam_w = ref.common.origin_to_world(R, scene["camera_mat"], world_mat, scene["scale_mat"])
    pix_w = ref.common.image_points_to_world(p, scene["camera_mat"], world_mat, scene["scale_mat"])
    rv = pix_w - cam_w
    rv = rv / rv.norm(2, 2).unsqueeze(-1)
    nerf.eval()
    with torch.no_grad():
        d_i = renderer.ray_marching(cam_w, rv, nerf, n_secant_steps=8, n_steps=[512, 513], rad=cfg["rendering"]["radius"])
    ph = renderer.phong_renderer(p, scene["camera_mat"], world_mat, scene["scale_mat"], 0)
    out["phong"] = {"pixels": p, "world_mat": world_mat, "camera_world": cam_w[0, 0].clone(), "ray_dirs": rv[0].clone(), "d_i": d_i,
                    "rgb": ph["rgb"].detach(), "rgb_surf": ph["rgb_surf"].detach()}
    print("geometry: %d of %d rays hit a surface, %d start occupied" % (int(torch.isfinite(d_i).sum() - (d_i == 0).sum()), R, int((d_i == 0).sum())))
    return out


def cfg2_scene_setup():
    """Inputs of the BASELINE.json cfg2 golden case: shared by this generator and tests/test_gpu_cfg2.py."""
    scene = syn.make_scene("cfg2")
    n = scene["n"]
    r0, t0 = syn.init_pose_params(n)
    scales = 1.0 + 0.01 * torch.arange(n, dtype=torch.float32).view(n, 1)
    shifts = 0.005 * torch.arange(n, dtype=torch.float32).view(n, 1)
    R = syn.SHAPES["cfg2"][3]
    return scene, r0, t0, scales, shifts, syn.ray_choice(scene["h"], scene["w"], R), syn.jitter(R, 128)


def _opt_clone(t):
    return t.clone() if t is not None else None


def cfg2_vectors(ref, cfg):
    """BASELINE.json cfg2 at full size through the UNMODIFIED reference Trainer.compute_loss + backward: 20-frame 375 x 1242
    scene, 4,096 injected rays x 128 samples, full loss set (rgb L1, depth l1 with learned scale/shift, Chamfer on 28,830-point
    clouds, reprojection), LearnPose.  ~25 GB and several minutes on the host."""
    import time
    scene, r0, t0, scales, shifts, ray_idx, noise = cfg2_scene_setup()
    n, h, w, R = scene["n"], scene["h"], scene["w"], ray_idx.numel()
    sd = syn.init_mlp_state(42)
    out = {"torch_version": torch.__version__, "cases": {}}
    for name, c in {"mid": dict(idx=5, ref=6), "last": dict(idx=n - 1, ref=n - 2)}.items():
        tcfg = dict(cfg["training"], n_training_points=R)
        nerf, renderer, net = build_ref_model(ref, cfg, sd)
        pose_net = ref.model.LearnPose(n, True, True, cfg, init_c2w=scene["poses_gt"].clone())
        dist_net = ref.model.Learn_Distortion(n, True, True, cfg)
        with torch.no_grad():
            pose_net.r.copy_(r0)
            pose_net.t.copy_(t0)
            dist_net.global_scales.copy_(scales)
            dist_net.global_shifts.copy_(shifts)
        trainer = ref.model.Trainer(net, None, tcfg, device=torch.device("cpu"), pose_param_net=pose_net, distortion_net=dist_net)
        # record what the unmodified loss functions see and return (the wrappers call the originals)
        seen = {}
        o_idx, o_pc = trainer.loss.comp_closest_pts_idx_with_split, trainer.loss.get_pc_loss

        def idx_rec(a, b, _o=o_idx):
            r = _o(a, b)
            seen.setdefault("nn_idx", []).append(r.to(torch.int32).clone())
            return r

        def pc_rec(X, Y, _o=o_pc):
            seen["X"], seen["Y"] = X.detach().clone(), Y.detach().clone()
            return _o(X, Y)

        trainer.loss.comp_closest_pts_idx_with_split, trainer.loss.get_pc_loss = idx_rec, pc_rec
        o_model_fwd = net.forward

        def net_rec(*a, _o=o_model_fwd, **k):
            r = _o(*a, **k)
            seen["rgb"], seen["depth_pred"], seen["depth_gt"] = r["rgb"].detach().clone(), r["depth_pred"].detach().clone(), r["depth_gt"].detach().clone()
            return r

        net.forward = net_rec
        batch = syn.make_batch(scene, c["idx"], c["ref"])
        t0_ = time.time()
        with injected_draws(noise=noise, perm=torch.cat([ray_idx, torch.zeros(h * w - R, dtype=torch.long)])):
            ld = trainer.compute_loss(batch, it=1, epoch=0, scheduling_start=tcfg["scheduling_start"], out_render_path="/tmp")
        ld["loss"].backward()
        small = ("fc_density.weight", "fc_density.bias", "fc_rgb.weight", "fc_rgb.bias", "rgb_layers.0.bias", "layers0.0.weight",
                 "layers1.6.bias")
        out["cases"][name] = {
            "case": c, "loss_dict": {k: v.detach().clone().reshape(-1) for k, v in ld.items()},
            "grad_r": pose_net.r.grad.clone(), "grad_t": pose_net.t.grad.clone(),
            "grad_scales": _opt_clone(dist_net.global_scales.grad), "grad_shifts": _opt_clone(dist_net.global_shifts.grad),
            "mlp_grad_stats": {k: flat_stats(v.grad) for k, v in nerf.named_parameters()},
            "mlp_grads": {k: v.grad.clone() for k, v in nerf.named_parameters() if k in small},
            "X": seen["X"][0].clone(), "Y": seen["Y"][0].clone(), "nn_idx_xy": seen["nn_idx"][0], "nn_idx_yx": seen["nn_idx"][1],
            "rgb": seen["rgb"], "depth_pred": seen["depth_pred"], "depth_gt": seen["depth_gt"],
        }
        print("cfg2", name, "%.0f s" % (time.time() - t0_), {k: float(v.reshape(-1)[0]) for k, v in ld.items()}, flush=True)
    out["setup"] = {"ray_idx_checksum": int(ray_idx.sum()), "noise_checksum": float(noise.double().sum()), "scales": scales, "shifts": shifts}
    return out


def cfg5_clouds():
    """BASELINE.json cfg5: (i) two 100,000-point clouds back-projected from a 250 x 400 depth crop in raster
    order, the second moved by a small rigid offset; (ii) two clouds drawn from U(-1,1)^3 (seeds 7 / 8)."""
    g = torch.Generator().manual_seed(7)
    hh, ww = 250, 400
    v, u = torch.meshgrid(torch.linspace(-0.75, 0.75, hh), torch.linspace(-1, 1, ww), indexing="ij")
    zz = 2.0 + 0.3 * torch.sin(3 * u) * torch.cos(2 * v)
    a = torch.stack([u * zz, v * zz, -zz], -1).reshape(-1, 3).contiguous()
    b = (a + torch.tensor([0.013, -0.007, 0.02]) + 1e-3 * torch.randn(a.shape, generator=g)).contiguous()
    ra = torch.rand(100000, 3, generator=torch.Generator().manual_seed(7)) * 2 - 1
    rb = torch.rand(100000, 3, generator=torch.Generator().manual_seed(8)) * 2 - 1
    return {"raster": (a, b), "random": (ra, rb)}


CFG5_SLICES = (0, 33333, 66666, 98976)   # starts of the four 1,024-source slices


def cfg5_vectors(ref, cfg):
    """Nearest-neighbour indices of the reference's comp_closest_pts_idx_with_split (model/losses.py:129-144) at 100,000 x
    100,000, evaluated on 1,024-source slices (the function's own kind of split; 1.2 GB per slice)."""
    L = ref.losses.Loss(cfg["training"])
    out = {"torch_version": torch.__version__, "slices": CFG5_SLICES, "slice_len": 1024}
    for name, (a, b) in cfg5_clouds().items():
        rec = {}
        for s0 in CFG5_SLICES:
            rec[s0] = L.comp_closest_pts_idx_with_split(a[s0:s0 + 1024].t().contiguous(), b.t().contiguous()).to(torch.int32)
        out[name] = rec
        print("cfg5", name, [int(v.sum()) for v in rec.values()], flush=True)
    return out


TRAJ_STEPS = 12
TRAJ_SMALL = ("fc_density.weight", "fc_density.bias", "fc_rgb.weight", "fc_rgb.bias", "layers0.0.bias", "layers1.6.bias", "rgb_layers.0.bias")


def trajectory_frames(n, t):
    """(frame, reference frame) of step t: cycles through the frames, the last one pairing with its predecessor."""
    idx = t % n
    return idx, (idx + 1 if idx < n - 1 else idx - 1)


def trajectory_vectors(ref, cfg):
    """TRAJ_STEPS consecutive calls of the UNMODIFIED reference's Trainer.train_step (model/training.py:68-100) with three
    torch.optim.Adam optimisers (train.py:59,100,118) on the tiny scene; ray choice and jitter of every step injected and
    recorded.  Steps 0-7 run at epoch 0 (l1 colour term, start weights), steps 8-9 in the middle of the annealing window
    (interpolated weights), steps 10-11 after it (l2 colour term, end weights).  Recorded: every loss term of every step and the parameters after the last one."""
    scene = syn.make_scene("tiny")
    n, h, w, R = scene["n"], scene["h"], scene["w"], syn.SHAPES["tiny"][3]
    sd = syn.init_mlp_state(42)
    r0, t0 = syn.init_pose_params(n)
    tcfg = dict(cfg["training"], n_training_points=R)
    nerf, renderer, net = build_ref_model(ref, cfg, sd)
    pose_net = ref.model.LearnPose(n, True, True, cfg, init_c2w=scene["poses_gt"].clone())
    dist_net = ref.model.Learn_Distortion(n, True, True, cfg)
    scales = 1.0 + 0.05 * torch.arange(n, dtype=torch.float32).view(n, 1)
    shifts = 0.01 * torch.arange(n, dtype=torch.float32).view(n, 1)
    with torch.no_grad():
        pose_net.r.copy_(r0)
        pose_net.t.copy_(t0)
        dist_net.global_scales.copy_(scales)
        dist_net.global_shifts.copy_(shifts)
    lrs = {"model": tcfg["learning_rate"], "pose": tcfg["pose_lr"], "distortion": tcfg["distortion_lr"]}
    trainer = ref.model.Trainer(net, torch.optim.Adam(net.parameters(), lr=lrs["model"]), tcfg, device=torch.device("cpu"),
                                optimizer_pose=torch.optim.Adam(pose_net.parameters(), lr=lrs["pose"]), pose_param_net=pose_net,
                                optimizer_distortion=torch.optim.Adam(dist_net.parameters(), lr=lrs["distortion"]), distortion_net=dist_net)
    g = torch.Generator().manual_seed(77)
    sched = tcfg["scheduling_start"]
    steps = []
    for t in range(TRAJ_STEPS):
        idx, ref_idx = trajectory_frames(n, t)
        epoch = 0 if t < 8 else (sched + tcfg["annealing_epochs"] // 2 if t < 10 else sched + tcfg["annealing_epochs"] + 1)
        perm = torch.randperm(h * w, generator=g)
        noise = torch.rand(1, R, 128, generator=g)
        with injected_draws(noise=noise, perm=perm):
            ld = trainer.train_step(syn.make_batch(scene, idx, ref_idx), it=t, epoch=epoch, scheduling_start=sched, render_path="/tmp")
        steps.append({"idx": idx, "ref": ref_idx, "epoch": epoch, "ray_idx": perm[:R].clone(), "noise": noise.reshape(R, 128).clone(),
                      "loss_dict": {k: v.detach().clone().reshape(-1) for k, v in ld.items()},
                      # parameters AFTER this step's optimiser updates (small tensors only)
                      "after": {"r": pose_net.r.detach().clone(), "t": pose_net.t.detach().clone(),
                                "scales": dist_net.global_scales.detach().clone(), "shifts": dist_net.global_shifts.detach().clone(),
                                **{k: v.detach().clone() for k, v in nerf.named_parameters() if k in TRAJ_SMALL}}})
        print("trajectory step", t, idx, ref_idx, epoch, {k: round(float(v.reshape(-1)[0]), 6) for k, v in ld.items()}, flush=True)
    return {"steps": steps, "lrs": lrs, "setup": {"scales": scales, "shifts": shifts, "r": r0, "t": t0},
            "final": {"r": pose_net.r.detach().clone(), "t": pose_net.t.detach().clone(),
                      "scales": dist_net.global_scales.detach().clone(), "shifts": dist_net.global_shifts.detach().clone(),
                      "mlp_stats": {k: flat_stats(v.detach()) for k, v in nerf.named_parameters()},
                      "mlp_small": {k: v.detach().clone() for k, v in nerf.named_parameters() if k in TRAJ_SMALL}}}


POSEFIT_STEPS = 8


def posefit_vectors(ref, cfg):
    """POSEFIT_STEPS calls of the UNMODIFIED reference's Trainer_pose.train_step (model/eval_pose_one_epoch.py:25-98: the test-time
    pose fit of evaluation/eval.py:122-137 -- MLP in eval mode, colour MSE, one Adam on the poses) on the tiny scene, poses
    started away from the truth; the ray choice of every step injected and recorded."""
    scene = syn.make_scene("tiny")
    n, h, w, R = scene["n"], scene["h"], scene["w"], syn.SHAPES["tiny"][3]
    sd = syn.init_mlp_state(42)
    r0, t0 = syn.init_pose_params(n)
    nerf, renderer, net = build_ref_model(ref, cfg, sd)
    pose_net = ref.model.LearnPose(n, True, True, cfg, init_c2w=scene["poses_gt"].clone())
    with torch.no_grad():
        pose_net.r.copy_(r0)
        pose_net.t.copy_(t0)
    lr = 1e-3
    trainer = ref.model.Trainer_pose(net, {"n_points": R, "type": "nope_nerf"}, device=torch.device("cpu"),
                                     optimizer_pose=torch.optim.Adam(pose_net.parameters(), lr=lr), pose_param_net=pose_net)
    g = torch.Generator().manual_seed(78)
    steps = []
    for t in range(POSEFIT_STEPS):
        idx = t % n
        batch = syn.make_batch(scene, idx, (idx + 1) % n)
        data = {"img": batch["img"], "img.idx": batch["img.idx"], "img.camera_mat": batch["img.camera_mat"], "img.scale_mat": batch["img.scale_mat"]}
        if t >= POSEFIT_STEPS // 2:
            data["img.depth"] = batch["img.dpt"]        # optional key (model/eval_pose_one_epoch.py:56): both forms
        perm = torch.randperm(h * w, generator=g)
        with injected_draws(perm=perm):
            ld = trainer.train_step(data)
        steps.append({"idx": idx, "ray_idx": perm[:R].clone(), "with_depth": "img.depth" in data, "loss": ld["loss"].detach().clone().reshape(-1),
                      "r": pose_net.r.detach().clone(), "t": pose_net.t.detach().clone()})
        print("posefit step", t, idx, float(ld["loss"]), flush=True)
    grads = {k: (v.grad is not None) for k, v in nerf.named_parameters()}
    return {"steps": steps, "lr": lr, "setup": {"r": r0, "t": t0}, "mlp_has_grad": grads}


def extract_vectors(ref, cfg):
    """The UNMODIFIED reference's Extract_Images.generate_images (model/extracting_images.py:39-131, the body of the loop in
    vis/render.py:104-109) on two frames of the tiny scene: the returned img / depth / disp arrays, the raw depth it saves as
    .npy, and the relative paths of everything it writes (imageio.imwrite and np.save are recorded instead of touching disk)."""
    import importlib
    import tempfile
    import numpy as np
    written = []
    sys.modules["imageio"].imwrite = lambda path, arr: written.append((path, np.array(arr)))
    ref_ex = importlib.import_module("model.extracting_images")
    scene = syn.make_scene("tiny")
    n, h, w = scene["n"], scene["h"], scene["w"]
    nerf, renderer, net = build_ref_model(ref, cfg, syn.init_mlp_state(42))
    r0, t0 = syn.init_pose_params(n)
    pose_net = ref.model.LearnPose(n, True, True, cfg, init_c2w=scene["poses_gt"].clone())
    with torch.no_grad():
        pose_net.r.copy_(r0)
        pose_net.t.copy_(t0)
        c2ws = torch.stack([pose_net(i) for i in range(n)])
    ex = ref_ex.Extract_Images(renderer, {"extract_images": {"resolution": [h, w]}}, use_learnt_poses=True, use_learnt_focal=False,
                               device=torch.device("cpu"), render_type="nope_nerf")
    out = {"c2ws": c2ws, "frames": {}}
    o_save = np.save
    with tempfile.TemporaryDirectory() as tmp:
        for i in (0, 2):
            del written[:]
            saved = []
            np.save = lambda path, arr: saved.append((path, np.array(arr)))
            try:
                data = {"img.camera_mat": scene["camera_mat"], "img.scale_mat": scene["scale_mat"], "img.idx": torch.tensor([i])}
                d = ex.generate_images(data, tmp, c2ws, None, 1000, False)
            finally:
                np.save = o_save
            out["frames"][i] = {"img": torch.from_numpy(np.ascontiguousarray(d["img"])), "depth": torch.from_numpy(np.ascontiguousarray(d["depth"])),
                                "disp": torch.from_numpy(np.ascontiguousarray(d["disp"])), "geo_is_none": d["geo"] is None,
                                "depth_raw": torch.from_numpy(saved[0][1]),
                                "paths": sorted(os.path.relpath(p, tmp) for p, _ in written + saved)}
            print("extract frame", i, out["frames"][i]["paths"], float(saved[0][1].mean()), flush=True)
    # Trainer.render_visdata (model/training.py:103-165; train.py:239): the preview of one frame with vis_geo on (the default)
    from PIL import Image
    tcfg = dict(cfg["training"], vis_geo=True)
    _, _, net_geo = build_ref_model(ref, cfg, geometry_state(syn.init_mlp_state(42)))     # a field WITH surfaces
    pose_vis = ref.model.LearnPose(n, True, True, cfg, init_c2w=scene["poses_gt"].clone())   # r = t = 0: the cameras of geometry.pt, which see surfaces
    trainer = ref.model.Trainer(net_geo, None, tcfg, device=torch.device("cpu"), pose_param_net=pose_vis)
    res = [12, 16]
    with tempfile.TemporaryDirectory() as tmp:
        del written[:]
        ret = trainer.render_visdata(syn.make_batch(scene, 1, 2), res, 1000, tmp)
        files = {f: torch.from_numpy(np.array(Image.open(os.path.join(tmp, f)))) for f in sorted(os.listdir(tmp))}
        for path, arr in written:
            files[os.path.relpath(path, tmp)] = torch.from_numpy(arr)
    out["visdata"] = {"resolution": res, "idx": 1, "returned": torch.from_numpy(np.ascontiguousarray(ret)), "files": files}
    print("visdata", {k: tuple(v.shape) for k, v in files.items()}, ret.shape, float(ret.mean()), flush=True)
    return out


def eval_lpips_stand_in(a, b, normalize=True):
    """Stands in for the LPIPS network evaluation/eval.py:163 builds (lpips is not installed): any callable with its signature."""
    return ((a - b).abs().mean() * (2.0 if normalize else 1.0)).reshape(1)


def eval_vectors(ref, cfg):
    """The UNMODIFIED reference's Eval_Images.eval_images (model/eval_images.py:47-219, the loop body of evaluation/eval.py:178-188)
    on two frames of the tiny scene: ground-truth depth at another resolution than the render (exercises the nearest-exact resize),
    sc = 0.5, a depth range that cuts both maps.  Returned dict and every file it writes are recorded."""
    import importlib
    import tempfile
    import numpy as np
    import cv2
    from oracle.ref_harness import REFERENCE_ROOT
    ref_harness._stub("skimage").metrics = None
    written = []
    sys.modules["imageio"].imwrite = lambda path, arr: written.append((path, np.array(arr)))
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        ref_ev = importlib.import_module("model.eval_images")
    finally:
        sys.path.remove(REFERENCE_ROOT)
    scene = syn.make_scene("tiny")
    n, h, w = scene["n"], scene["h"], scene["w"]
    nerf, renderer, net = build_ref_model(ref, cfg, syn.init_mlp_state(42))
    c2ws = scene["poses_gt"].clone()
    ev = ref_ev.Eval_Images(renderer, {"extract_images": {"resolution": [h, w]}}, use_learnt_poses=True, use_learnt_focal=False,
                            device=torch.device("cpu"), render_type="nope_nerf", c2ws=c2ws, img_list=None)
    out = {"c2ws": c2ws, "frames": {}, "sc": 0.5, "min_depth": 0.03, "max_depth": 0.036}
    g = torch.Generator().manual_seed(5)
    for i in (1, 3):
        batch = syn.make_batch(scene, i, (i + 1) % n)
        gt_depth = 0.03 + 0.01 * torch.rand(1, h // 2, w // 2, generator=g)          # (1, 12, 16): half the render resolution
        data = {"img": batch["img"], "img.idx": batch["img.idx"], "img.camera_mat": batch["img.camera_mat"], "img.scale_mat": batch["img.scale_mat"],
                "img.gt_depths": gt_depth}
        with tempfile.TemporaryDirectory() as tmp:
            del written[:]
            d = ev.eval_images(data, tmp, None, eval_lpips_stand_in, logger=None, min_depth=out["min_depth"], max_depth=out["max_depth"], sc=out["sc"],
                               show_errors=False)
            files = {os.path.relpath(p, tmp): torch.from_numpy(a) for p, a in written}
            for r, _, fs in os.walk(tmp):
                for f in fs:
                    files[os.path.relpath(os.path.join(r, f), tmp)] = torch.from_numpy(cv2.imread(os.path.join(r, f), cv2.IMREAD_UNCHANGED))
        out["frames"][i] = {"gt_depth": gt_depth, "files": files,
                            "ret": {k: (torch.from_numpy(np.ascontiguousarray(v)) if isinstance(v, np.ndarray) else float(v)) for k, v in d.items()}}
        print("eval frame", i, {k: (tuple(v.shape) if isinstance(v, np.ndarray) else round(float(v), 5)) for k, v in d.items()}, sorted(files), flush=True)
    return out


def main():
    os.makedirs(OUT, exist_ok=True)
    if "--evalimg" in sys.argv:
        ref = ref_harness.load_reference()
        torch.set_num_threads(max(1, os.cpu_count() or 1))
        torch.save(eval_vectors(ref, syn.default_cfg()), os.path.join(OUT, "evalimg.pt"))
        return
    if "--trajectory" in sys.argv or "--posefit" in sys.argv or "--extract" in sys.argv:
        ref = ref_harness.load_reference()
        torch.set_num_threads(max(1, os.cpu_count() or 1))
        if "--extract" in sys.argv:
            torch.save(extract_vectors(ref, syn.default_cfg()), os.path.join(OUT, "extract.pt"))
        if "--trajectory" in sys.argv:
            torch.save(trajectory_vectors(ref, syn.default_cfg()), os.path.join(OUT, "trajectory.pt"))
        if "--posefit" in sys.argv:
            torch.save(posefit_vectors(ref, syn.default_cfg()), os.path.join(OUT, "posefit.pt"))
        return
    if "--cfg2" in sys.argv or "--cfg5" in sys.argv:
        ref = ref_harness.load_reference()
        torch.set_num_threads(max(1, os.cpu_count() or 1))
        if "--cfg5" in sys.argv:
            torch.save(cfg5_vectors(ref, syn.default_cfg()), os.path.join(OUT, "cfg5_chamfer.pt"))
        if "--cfg2" in sys.argv:
            torch.save(cfg2_vectors(ref, syn.default_cfg()), os.path.join(OUT, "cfg2_step.pt"))
        return
    if "--extra-step-cases" in sys.argv:
        # adds the later step cases to render_step.pt, leaving every existing entry byte-identical
        os.environ["NNB_GOLDEN_EXTRA_CASES"] = "1"
        ref = ref_harness.load_reference()
        torch.set_num_threads(max(1, os.cpu_count() or 1))
        extra = render_and_step_vectors(ref, syn.default_cfg())["step"]
        path = os.path.join(OUT, "render_step.pt")
        g = torch.load(path)
        g["step"].update(extra)
        torch.save(g, path)
        return
    if "--geometry-only" in sys.argv:
        ref = ref_harness.load_reference()
        torch.set_num_threads(max(1, os.cpu_count() or 1))
        torch.save(geometry_vectors(ref, syn.default_cfg()), os.path.join(OUT, "geometry.pt"))
        return
    if "--variants-only" in sys.argv:
        ref = ref_harness.load_reference()
        torch.set_num_threads(max(1, os.cpu_count() or 1))
        torch.save(variant_vectors(ref, syn.default_cfg()), os.path.join(OUT, "render_variants.pt"))
        return
    ref = ref_harness.load_reference()
    cfg = syn.default_cfg()
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    torch.save(stage_vectors(ref, cfg), os.path.join(OUT, "stages.pt"))
    torch.save(loss_vectors(ref, cfg), os.path.join(OUT, "losses.pt"))
    torch.save(render_and_step_vectors(ref, cfg), os.path.join(OUT, "render_step.pt"))
    torch.save(variant_vectors(ref, cfg), os.path.join(OUT, "render_variants.pt"))
    torch.save(geometry_vectors(ref, cfg), os.path.join(OUT, "geometry.pt"))
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))


if __name__ == "__main__":
    main()
