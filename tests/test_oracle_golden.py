"""Pins the CPU oracle (oracle/nope_oracle.py) against fixtures produced by the reference itself
(oracle/make_golden.py).  CPU only.  Tolerances: fp32 paths 1e-5 relative (north_star)."""
import pytest
import torch

from oracle import nope_oracle as O
import synthetic_scene as syn

RTOL = 1e-5


def close(a, b, rtol=RTOL, atol=1e-6):
    torch.testing.assert_close(a, b, rtol=rtol, atol=atol)


def test_pose_exp_and_grad(golden):
    g = golden("stages")["pose"]
    for i in range(g["r"].shape[0]):
        r = g["r"][i].clone().requires_grad_(True)
        t = g["t"][i].clone().requires_grad_(True)
        close(O.pose_c2w(r, t), g["c2w"][i])
        m = O.pose_c2w(r, t, g["init"][i])
        close(m, g["c2w_init"][i])
        (m * g["probe"]).sum().backward()
        close(r.grad, g["grad_r"][i], atol=1e-5)
        close(t.grad, g["grad_t"][i])
    # Exp(0) is exactly the identity
    assert torch.equal(O.exp_so3(torch.zeros(3)), torch.eye(3))


def test_distortion(golden):
    g = golden("stages")["distortion"]
    for i in range(4):
        s, sh = O.distortion(g["scales"], g["shifts"], i)
        assert abs(float(s) - g["out"][i][0]) < 1e-7 and abs(float(sh) - g["out"][i][1]) < 1e-7
    assert g["out"][1][0] == 0.01 or abs(g["out"][1][0] - 0.01) < 1e-8   # clamped
    assert g["out"][3][0] == 1.0                                         # fix_scaleN


def test_pixel_grid(golden):
    g = golden("stages")["pixel_grid"]
    loc, sc = O.pixel_grid(g["h"], g["w"])
    assert torch.equal(loc, g["loc"])
    assert torch.equal(sc, g["scaled"])


def test_posenc(golden):
    g = golden("stages")["posenc"]
    assert torch.equal(O.posenc(g["x"], 10), g["enc10"])
    assert torch.equal(O.posenc(g["x"], 4), g["enc4"])


def test_composite(golden):
    g = golden("stages")["composite"]
    a = g["alpha"].clone().requires_grad_(True)
    c = g["rgb_s"].clone().requires_grad_(True)
    rgb, dist, w, _ = O.composite(a, c, g["z"])
    close(w, g["weights"])
    close(rgb, g["rgb"])
    close(dist, g["dist"])
    ((rgb * g["probe_rgb"]).sum() + (dist * g["probe_dist"]).sum()).backward()
    close(a.grad, g["grad_alpha"], atol=1e-5)
    close(c.grad, g["grad_rgb_s"])


def test_losses(golden):
    g = golden("losses")
    close(O.rgb_loss(g["rgb"]["pred"], g["rgb"]["gt"], "l1"), g["rgb"]["l1"])
    close(O.rgb_loss(g["rgb"]["pred"], g["rgb"]["gt"], "l2"), g["rgb"]["l2"])
    for n in (50, 51):
        d = g["depth_%d" % n]
        p = d["pred"].clone().requires_grad_(True)
        q = d["gt"].clone().requires_grad_(True)
        close(O.depth_l1(p, q), d["l1"])
        inv = O.depth_invariant(p, q)
        close(inv, d["inv"])
        inv.backward()
        close(p.grad, d["inv_grad_pred"])
        close(q.grad, d["inv_grad_gt"])
    s = g["rgb_s"]
    close(O.photometric_loss(s["rgb1"], s["rgb2"], s["valid"]), s["loss"])
    assert float(O.photometric_loss(s["rgb1"], s["rgb2"], torch.zeros_like(s["valid"]))) == 0.0


def test_chamfer(golden):
    c = golden("losses")["chamfer"]
    assert torch.equal(O.nearest_idx(c["X"], c["Y"]), c["idx_xy"])
    assert torch.equal(O.nearest_idx(c["Y"], c["X"], slice_len=100), c["idx_yx"])
    assert torch.equal(O.nearest_idx(c["lat_src"], c["lat_dst"]), c["lat_idx"])  # exact ties -> first index
    X = c["X"].clone().requires_grad_(True)
    Y = c["Y"].clone().requires_grad_(True)
    loss = O.chamfer(X.t().unsqueeze(0), Y.t().unsqueeze(0))
    close(loss, c["loss"])
    loss.backward()
    close(X.grad, c["grad_X"])
    close(Y.grad, c["grad_Y"])
    assert torch.isfinite(X.grad).all()  # coincident pair: zero subgradient, not NaN


def _params(requires_grad=False):
    sd = syn.init_mlp_state(42)
    return {k: v.clone().requires_grad_(requires_grad) for k, v in sd.items()}


def test_mlp_weights_reproducible(golden):
    g = golden("render_step")
    sd = syn.init_mlp_state(42)
    for k, v in sd.items():
        close(torch.stack([v.sum(), v.abs().sum(), (v * v).sum()]), g["mlp_checksum"][k])


def test_mlp_forward(golden):
    g = golden("render_step")["mlp"]
    with torch.no_grad():
        rgb, alpha = O.mlp_forward(_params(), g["pts"], g["viewdir"])
    close(rgb, g["rgb"])
    close(alpha, g["alpha"])


def test_render_train_and_eval(golden):
    g = golden("render_step")["render"]
    cfg = syn.default_cfg()
    scene = syn.make_scene("tiny")
    P = _params()
    with torch.no_grad():
        tr = O.render(P, g["pixels"], g["depth"], scene["camera_mat"], g["world_mat"], scene["scale_mat"],
                      cfg["rendering"], noise=g["noise"][0], eval_=False)
        ev = O.render(P, g["pixels"], g["depth"], scene["camera_mat"], g["world_mat"], scene["scale_mat"],
                      cfg["rendering"], noise=None, eval_=True)
    for mine, ref in ((tr, g["train"]), (ev, g["eval"])):
        for k in ("rgb", "z_vals", "depth_pred", "depth_gt", "alpha"):
            assert mine[k].shape == ref[k].shape, k
            close(mine[k], ref[k], atol=2e-6)
    assert tr["depth_pred"].shape[0] == g["pixels"].shape[1] - 1   # one masked ray


def test_train_step_loss_and_grads(golden):
    g = golden("render_step")
    cfg = syn.default_cfg()
    scene = syn.make_scene("tiny")
    R = syn.SHAPES["tiny"][3]
    ray_idx, noise = g["render"]["ray_idx"], g["render"]["noise"]
    for name, rec in g["step"].items():
        c = rec["case"]
        cfg["training"]["depth_loss_type"] = c["depth_loss_type"]
        cfg["training"]["shift_first"] = c.get("shift_first", False)
        P = _params(True)
        pose = {"r": g["step_setup"]["r"].clone().requires_grad_(True), "t": g["step_setup"]["t"].clone().requires_grad_(True),
                "init_c2w": scene["poses_gt"]}
        dist = {"global_scales": g["step_setup"]["scales"].clone().requires_grad_(True),
                "global_shifts": g["step_setup"]["shifts"].clone().requires_grad_(True)}
        batch = syn.make_batch(scene, c["idx"], c["ref"])
        ld, _, _ = O.train_loss(batch, P, pose, dist, cfg, ray_idx, noise, epoch=c["epoch"])
        for k in O.LOSS_KEYS + ("scale", "shift"):
            close(ld[k].reshape(-1).float(), rec["loss_dict"][k].float(), atol=2e-6)
        ld["loss"].backward()
        close(pose["r"].grad, rec["grad_r"], rtol=1e-4, atol=1e-5)
        close(pose["t"].grad, rec["grad_t"], rtol=1e-4, atol=1e-5)
        if rec["grad_scales"] is not None:
            close(dist["global_scales"].grad, rec["grad_scales"], rtol=1e-4, atol=1e-5)
            close(dist["global_shifts"].grad, rec["grad_shifts"], rtol=1e-4, atol=1e-5)
        for k, st in rec["mlp_grad_stats"].items():
            gr = P[k].grad
            close(torch.stack([gr.sum(), gr.abs().sum(), (gr * gr).sum()]), st, rtol=1e-4, atol=1e-6)
        if "mlp_grads" in rec:
            for k, v in rec["mlp_grads"].items():
                close(P[k].grad, v, rtol=1e-4, atol=1e-6)


def _variant_cfg(over):
    import copy
    cfg = copy.deepcopy(syn.default_cfg())
    for sec, kv in over.items():
        cfg[sec].update(kv)
    return cfg


@pytest.mark.parametrize("name", ["llff", "dist_alpha_uniform", "white_relu_nodir"])
def test_render_variants(golden, name):
    """Non-default renderer branches against the reference's own Renderer: NDC sampling + dist_alpha
    (configs/LLFF/fern.yaml), dist_alpha with uniform sampling, white background + relu density + use_ray_dir False."""
    g = golden("render_variants")
    i, v = g["inputs"], g[name]
    cfg = _variant_cfg(v["override"])
    scene = syn.make_scene("tiny")
    P = _params(True)
    wm = i["world_mat"].clone().requires_grad_(True)
    cm = i["camera_mat"].clone().requires_grad_(True)
    tr = O.render(P, i["pixels"], i["depth"], cm, wm, scene["scale_mat"], cfg["rendering"], noise=i["noise"][0], eval_=False,
                  occ_activation=cfg["model"]["occ_activation"])
    for k in ("rgb", "z_vals", "depth_pred", "depth_gt", "alpha"):
        assert tr[k].shape == v["train"][k].shape, k
        close(tr[k].detach(), v["train"][k], atol=2e-6)
    ((tr["rgb"] * i["probe_rgb"]).sum() + (tr["depth_pred"] * i["probe_dist"][:tr["depth_pred"].shape[0]]).sum()).backward()
    close(wm.grad, v["grad_world_mat"], rtol=1e-3, atol=1e-6)
    close(cm.grad, v["grad_camera_mat"], rtol=1e-3, atol=1e-6)
    for k, st in v["mlp_grad_stats"].items():
        gr = P[k].grad
        close(torch.stack([gr.sum(), gr.abs().sum(), (gr * gr).sum()]), st, rtol=1e-4, atol=1e-6)
    with torch.no_grad():
        ev = O.render(_params(), i["pixels"], i["depth"], i["camera_mat"], i["world_mat"], scene["scale_mat"], cfg["rendering"],
                      noise=None, eval_=True, occ_activation=cfg["model"]["occ_activation"])
    for k in ("rgb", "z_vals", "depth_pred", "depth_gt", "alpha"):
        close(ev[k], v["eval"][k], atol=2e-6)


def _geometry_params():
    sd = syn.init_mlp_state(42)
    sd = {k: v.clone() for k, v in sd.items()}
    sd["fc_density.weight"] = sd["fc_density.weight"] * 200.0     # oracle/make_golden.py::geometry_state
    sd["fc_density.bias"] = torch.tensor([0.86])
    return sd


def test_geometry_path(golden):
    """SURVEY.md 8f rank 3 against the reference's own gradient / ray_marching / phong_renderer."""
    g = golden("geometry")
    P = _geometry_params()
    gr = g["gradient"]
    close(O.density_gradient(P, gr["pts"]), gr["grad"], rtol=1e-4, atol=1e-5)
    with torch.no_grad():
        close(O.occupancy(P, gr["pts"]), gr["occ"], atol=2e-6)
    ph = g["phong"]
    rgb, rgb_surf, d = O.phong(P, ph["camera_world"], ph["ray_dirs"], syn.default_cfg()["rendering"]["radius"])
    close(d.unsqueeze(0), ph["d_i"], rtol=1e-4, atol=1e-5)
    close(rgb.unsqueeze(0), ph["rgb"], rtol=1e-3, atol=1e-4)
    close(rgb_surf.unsqueeze(0), ph["rgb_surf"], rtol=1e-3, atol=1e-4)
