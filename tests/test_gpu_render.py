"""GPU parity, stages 2+3 (sampling, positional encoding, MLP, compositing) against the reference's golden
outputs and the CPU oracle.  fp32 path tolerance: the north_star's 1e-5 relative is met per layer; through 12
layers + a 2^9 positional-encoding gain the end-to-end figure used here is 1e-4 relative / 2e-5 absolute
(measured errors are printed by the tests)."""
import pytest
import torch

from util import assert_close, dev, max_err, syn

pytestmark = pytest.mark.gpu


def _params_dev():
    sd = syn.init_mlp_state(42)
    from model import _cabi
    return sd, [sd[k].to(dev()) for k in _cabi.PARAM_ORDER]


def test_composite_vs_reference(golden):
    from model import _ops
    g = golden("stages")["composite"]
    a = g["alpha"].to(dev()).requires_grad_(True)
    c = g["rgb_s"].to(dev()).requires_grad_(True)
    rgb, dist, w = _ops.Composite.apply(a, c, g["z"].to(dev()), False, False)
    assert_close(w, g["weights"], 1e-5, 1e-7, "weights")
    assert_close(rgb, g["rgb"], 1e-5, 1e-6, "rgb")
    assert_close(dist, g["dist"], 1e-5, 1e-6, "dist")
    ((rgb * g["probe_rgb"].to(dev())).sum() + (dist * g["probe_dist"].to(dev())).sum()).backward()
    assert_close(a.grad, g["grad_alpha"], 1e-4, 1e-5, "d alpha")
    assert_close(c.grad, g["grad_rgb_s"], 1e-5, 1e-6, "d rgb_s")


def test_composite_dist_alpha_and_white_vs_oracle():
    from model import _ops
    from oracle import nope_oracle as O
    g = torch.Generator().manual_seed(4)
    sig = (torch.rand(33, 128, generator=g) * 3).requires_grad_(True)
    c = torch.rand(33, 128, 3, generator=g).requires_grad_(True)
    z = torch.sort(torch.rand(33, 128, generator=g) * 10)[0]
    rgb, dist, w, _ = O.composite(sig, c, z, dist_alpha=True, white_background=True)
    p1, p2 = torch.randn(33, 3, generator=g), torch.randn(33, generator=g)
    ((rgb * p1).sum() + (dist * p2).sum()).backward()
    sd, cd = sig.detach().to(dev()).requires_grad_(True), c.detach().to(dev()).requires_grad_(True)
    rgb2, dist2, w2 = _ops.Composite.apply(sd, cd, z.to(dev()), True, True)
    assert_close(w2, w, 1e-5, 1e-7, "weights")
    assert_close(rgb2, rgb, 1e-5, 1e-6, "rgb")
    ((rgb2 * p1.to(dev())).sum() + (dist2 * p2.to(dev())).sum()).backward()
    assert_close(sd.grad, sig.grad, 1e-4, 1e-5, "d sigma")
    assert_close(cd.grad, c.grad, 1e-5, 1e-6, "d rgb_s")


def test_mlp_points_vs_reference(golden):
    from model import _ops
    g = golden("render_step")["mlp"]
    _, params = _params_dev()
    rgb, alpha = _ops.mlp_points(params, g["pts"].to(dev()), g["viewdir"].to(dev()), True, False)
    print("mlp points max err rgb %.3e alpha %.3e" % (max_err(rgb, g["rgb"]), max_err(alpha, g["alpha"])))
    assert_close(rgb, g["rgb"], 1e-4, 2e-5, "rgb")
    assert_close(alpha, g["alpha"], 1e-4, 2e-5, "alpha")


@pytest.mark.parametrize("mode", ["train", "eval"])
def test_renderer_vs_reference(golden, mode):
    """Renderer.forward through the drop-in module, same inputs as the reference run that made the fixture."""
    from util import build_models
    g = golden("render_step")["render"]
    cfg = syn.default_cfg()
    scene = syn.make_scene("tiny")
    _, renderer, _, _, _ = build_models(cfg, scene["n"])
    args = (g["pixels"].to(dev()), g["depth"].to(dev()), scene["camera_mat"].to(dev()), g["world_mat"].to(dev()),
            scene["scale_mat"].to(dev()), "nope_nerf")
    with torch.no_grad():
        if mode == "train":
            out = renderer(*args, add_noise=True, eval_=False, noise=g["noise"].to(dev()))
        else:
            out = renderer(*args, add_noise=False, eval_=True)
    ref = g[mode]
    for k in ("rgb", "z_vals", "depth_pred", "depth_gt", "alpha"):
        print(mode, k, "max err %.3e" % max_err(out[k], ref[k]))
    assert torch.equal(out["z_vals"].cpu(), ref["z_vals"]) or max_err(out["z_vals"], ref["z_vals"]) < 2e-6
    assert out["depth_pred"].shape == ref["depth_pred"].shape
    assert_close(out["depth_gt"], ref["depth_gt"], 1e-5, 1e-6, "depth_gt")
    # end-to-end through ray generation: a 1-ulp change in a ray direction is a 2^9-ulp phase change in the top
    # positional-encoding band, so the fp32 figure is ~1e-4 here (stage-isolated tests below are tighter)
    assert_close(out["rgb"], ref["rgb"], 1e-3, 1e-4, "rgb")
    assert_close(out["alpha"], ref["alpha"], 1e-3, 1e-4, "alpha")
    assert_close(out["depth_pred"], ref["depth_pred"], 1e-3, 1e-4, "depth_pred")


def _oracle_render(requires_grad):
    from oracle import nope_oracle as O
    cfg = syn.default_cfg()
    scene = syn.make_scene("tiny")
    R = 64
    idx = syn.ray_choice(scene["h"], scene["w"], R)
    pixels = O.pixel_grid(scene["h"], scene["w"])[1][:, idx]
    depth = scene["depths"][1].reshape(1, -1, 1)[:, idx]
    r0, t0 = syn.init_pose_params(scene["n"])
    world = torch.inverse(O.pose_c2w(r0[1], t0[1], scene["poses_gt"][1])).unsqueeze(0)
    o, d, _, _, _ = O.ray_setup(pixels, depth, scene["camera_mat"], world, scene["scale_mat"])
    sd = syn.init_mlp_state(42)
    P = {k: v.clone().requires_grad_(requires_grad) for k, v in sd.items()}
    noise = syn.jitter(R)[0]
    return cfg, o.detach(), d.detach(), noise, P


def test_render_stage_forward_isolated():
    """Same origin/dirs bits on both sides: isolates sampling + encoding + MLP + compositing."""
    from model import _cabi, _ops
    from oracle import nope_oracle as O
    cfg, o, d, noise, P = _oracle_render(False)
    R, S = d.shape[0], 128
    with torch.no_grad():
        z = O.stratified_z(R, S, 0.01, 10.0, noise)
        pts = (o.reshape(1, 1, 3) + d.unsqueeze(1) * z.unsqueeze(-1)).reshape(-1, 3)
        view = (-d).unsqueeze(1).expand(R, S, 3).reshape(-1, 3)
        rgb_s, alpha = O.mlp_forward(P, pts, view)
        rgb, dist, w, _ = O.composite(alpha.reshape(R, S), rgb_s.reshape(R, S, 3), z)
    rc = _ops.make_cfg(R, S, 0.01, 10.0, 1, 0, 0, 1, _cabi.PREC_FP32)
    params = [P[k].to(dev()) for k in _cabi.PARAM_ORDER]
    rgb2, dist2, w2, a2, z2 = _ops.render_nograd(rc, params, None, o.to(dev()), d.to(dev()), noise.to(dev()))
    assert torch.equal(z2.cpu(), z), "sample depths must be bit-exact"
    print("isolated render max err: rgb %.3e dist %.3e alpha %.3e" % (max_err(rgb2, rgb), max_err(dist2, dist), max_err(a2, alpha.reshape(R, S))))
    assert_close(a2, alpha.reshape(R, S), 1e-4, 2e-5, "alpha")
    assert_close(w2, w, 1e-4, 2e-5, "weights")
    assert_close(rgb2, rgb, 1e-4, 2e-5, "rgb")
    assert_close(dist2, dist, 1e-4, 2e-5, "dist")


def test_render_stage_backward_isolated():
    from model import _cabi, _ops
    from oracle import nope_oracle as O
    cfg, o, d, noise, P = _oracle_render(True)
    R, S = d.shape[0], 128
    oc, dc = o.clone().requires_grad_(True), d.clone().requires_grad_(True)
    z = O.stratified_z(R, S, 0.01, 10.0, noise)
    pts = (oc.reshape(1, 1, 3) + dc.unsqueeze(1) * z.unsqueeze(-1)).reshape(-1, 3)
    view = (-dc).unsqueeze(1).expand(R, S, 3).reshape(-1, 3)
    rgb_s, alpha = O.mlp_forward(P, pts, view)
    rgb, dist, w, _ = O.composite(alpha.reshape(R, S), rgb_s.reshape(R, S, 3), z)
    g = torch.Generator().manual_seed(8)
    p_rgb, p_dist = torch.randn(R, 3, generator=g) / R, torch.randn(R, generator=g) / R
    ((rgb * p_rgb).sum() + (dist * p_dist).sum()).backward()

    rc = _ops.make_cfg(R, S, 0.01, 10.0, 1, 0, 0, 1, _cabi.PREC_FP32)
    params = [P[k].detach().to(dev()).requires_grad_(True) for k in _cabi.PARAM_ORDER]
    od, dd = o.to(dev()).requires_grad_(True), d.to(dev()).requires_grad_(True)
    rgb2, dist2, w2, a2, z2 = _ops.RenderRays.apply(rc, None, od, dd, noise.to(dev()), *params)
    ((rgb2 * p_rgb.to(dev())).sum() + (dist2 * p_dist.to(dev())).sum()).backward()
    worst = 0.0
    for k, p in zip(_cabi.PARAM_ORDER, params):
        ref = P[k].grad
        e = float((p.grad.cpu() - ref).abs().max() / ref.abs().max().clamp_min(1e-12))
        worst = max(worst, e)
        print("  grad %-24s scale-relative err %.3e" % (k, e))
    print("render bwd: worst scale-relative weight-grad err %.3e" % worst)
    assert worst < 2e-3
    assert_close(od.grad, oc.grad, 2e-3, 1e-6 + 2e-3 * float(oc.grad.abs().max()), "d origin")
    assert_close(dd.grad, dc.grad, 2e-3, 1e-6 + 2e-3 * float(dc.grad.abs().max()), "d dirs")


def _variant_cfg(over):
    import copy
    cfg = copy.deepcopy(syn.default_cfg())
    for sec, kv in over.items():
        cfg[sec].update(kv)
    return cfg


@pytest.mark.parametrize("precision", ["fp32", "bf16"])
@pytest.mark.parametrize("name", ["llff", "dist_alpha_uniform", "white_relu_nodir"])
def test_renderer_variants_vs_reference(golden, name, precision):
    """Non-default renderer branches through the drop-in Renderer against fixtures made by the
    reference's Renderer: NDC sampling + dist_alpha (configs/LLFF/fern.yaml), dist_alpha with uniform sampling, white
    background + relu density + use_ray_dir False.  fp32 path: 1e-3 relative end to end (2^9 positional-encoding gain);
    bf16 tensor-core path: 1e-2 abs on rgb/alpha (north_star), gradients 0.15 relative L2 at 48 rays (DESIGN.md sections 2 and 6)."""
    from util import build_models
    g = golden("render_variants")
    i, v = g["inputs"], g[name]
    cfg = _variant_cfg(v["override"])
    scene = syn.make_scene("tiny")
    nerf, renderer, _, _, _ = build_models(cfg, scene["n"], precision=precision)
    wm = i["world_mat"].to(dev()).requires_grad_(True)
    cm = i["camera_mat"].to(dev()).requires_grad_(True)
    out = renderer(i["pixels"].to(dev()), i["depth"].to(dev()), cm, wm, scene["scale_mat"].to(dev()), "nope_nerf",
                   add_noise=True, eval_=False, noise=i["noise"].to(dev()))
    ref = v["train"]
    for k in ("rgb", "z_vals", "depth_pred", "depth_gt", "alpha"):
        print(name, precision, k, "max err %.3e" % max_err(out[k], ref[k]))
        assert out[k].shape == ref[k].shape, k
    assert max_err(out["z_vals"], ref["z_vals"]) < 2e-6
    assert_close(out["depth_gt"], ref["depth_gt"], 1e-5, 1e-6, "depth_gt")
    far = cfg["rendering"]["depth_range"][1]
    if precision == "fp32":
        assert_close(out["rgb"], ref["rgb"], 1e-3, 1e-4, "rgb")
        assert_close(out["alpha"], ref["alpha"], 1e-3, 1e-4, "alpha")
        assert_close(out["depth_pred"], ref["depth_pred"], 1e-3, 1e-4 * far, "depth_pred")
    else:
        assert max_err(out["rgb"], ref["rgb"]) <= 1e-2 and max_err(out["alpha"], ref["alpha"]) <= 1e-2
        assert max_err(out["depth_pred"], ref["depth_pred"]) <= 1e-2 * far
    n = out["depth_pred"].shape[0]
    ((out["rgb"] * i["probe_rgb"].to(dev())).sum() + (out["depth_pred"] * i["probe_dist"][:n].to(dev())).sum()).backward()

    def rel_l2(a, b):
        return float((a.detach().cpu() - b).norm() / b.norm().clamp_min(1e-20))
    # fp32: forward agrees to ~1e-7; the first layer's weight gradient is a cancelling sum over 6,144 samples of
    # 2^9-gain encodings, so fp32 summation order alone moves it by ~3e-3 relative L2 (measured 2.9e-3 on the NDC case)
    tol = 5e-3 if precision == "fp32" else 0.15      # measured on B200 (48 rays): fp32 <= 2.9e-3, bf16 <= 0.10
    errs = {"world_mat": rel_l2(wm.grad, v["grad_world_mat"]), "camera_mat": rel_l2(cm.grad, v["grad_camera_mat"])}
    for k, gr in v["mlp_grads"].items():
        errs[k] = rel_l2(dict(nerf.named_parameters())[k].grad, gr)
    print(name, precision, "gradient rel-L2:", {k: "%.2e" % e for k, e in errs.items()})
    assert max(errs.values()) <= tol, errs
    # eval branch (no jitter, distances divided by the ray norm)
    with torch.no_grad():
        ev = renderer(i["pixels"].to(dev()), i["depth"].to(dev()), i["camera_mat"].to(dev()), i["world_mat"].to(dev()),
                      scene["scale_mat"].to(dev()), "nope_nerf", add_noise=False, eval_=True)
    etol = 1e-3 if precision == "fp32" else 1e-2
    assert max_err(ev["rgb"], v["eval"]["rgb"]) <= etol
    assert max_err(ev["depth_pred"], v["eval"]["depth_pred"]) <= etol * far
    assert_close(ev["depth_gt"], v["eval"]["depth_gt"], 1e-5, 1e-6, "eval depth_gt")


@pytest.mark.parametrize("dist_alpha,white", [(False, False), (True, False), (False, True)])
def test_composite_vector_path_is_bit_identical_to_the_generic_kernels(dist_alpha, white):
    """S = 128 with 16-byte aligned tensors takes the float4 kernels (composite_fwd128 / bwd128); the same values through views
    that start 4 bytes into their storage take the generic kernels.  Same arithmetic in the same order: outputs and gradients
    must agree bit for bit."""
    from model import _cabi
    from model._cabi import call, ptr, stream
    g = torch.Generator().manual_seed(17)
    R, S = 777, 128
    def pair(*shape, scale=1.0):
        t = (torch.rand(*shape, generator=g) * scale).to(dev())
        pad = torch.empty(t.numel() + 1, device=dev())
        mis = pad[1:].view(*shape)          # 4 bytes into the allocation: not 16-byte aligned
        mis.copy_(t)
        assert t.data_ptr() % 16 == 0 and mis.data_ptr() % 16 == 4
        return t, mis
    alpha, alpha_m = pair(R, S, scale=0.3)
    rgb_s, rgb_s_m = pair(R, S, 3)
    z0 = torch.sort(torch.rand(R, S, generator=g) * 10)[0].to(dev())
    z, z_m = z0.clone(), torch.empty(R * S + 1, device=dev())[1:].view(R, S)
    z_m.copy_(z0)
    g_rgb, g_dist = torch.randn(R, 3, generator=g).to(dev()), torch.randn(R, generator=g).to(dev())
    g_w, g_w_m = pair(R, S)
    outs = []
    for a_, c_, z_, gw_, aligned in ((alpha, rgb_s, z, g_w, True), (alpha_m, rgb_s_m, z_m, g_w_m, False)):
        rgb, dist = torch.empty(R, 3, device=dev()), torch.empty(R, device=dev())
        if aligned:
            w, ao, da, dc = (torch.empty(R, S, device=dev()), torch.empty(R, S, device=dev()), torch.empty(R, S, device=dev()),
                             torch.empty(R, S, 3, device=dev()))
        else:
            w, ao, da = [torch.empty(R * S + 1, device=dev())[1:].view(R, S) for _ in range(3)]
            dc = torch.empty(R * S * 3 + 1, device=dev())[1:].view(R, S, 3)
        call("nnb_composite_fwd", ptr(a_), ptr(c_), ptr(z_), R, S, int(dist_alpha), int(white), ptr(rgb), ptr(dist), ptr(w), ptr(ao), stream())
        call("nnb_composite_bwd", ptr(a_), ptr(c_), ptr(z_), ptr(w), R, S, int(dist_alpha), int(white), ptr(g_rgb), ptr(g_dist), ptr(gw_),
             ptr(da), ptr(dc), stream())
        torch.cuda.synchronize()
        outs.append([t.clone() for t in (rgb, dist, w, ao, da, dc)])
    for name, a, b in zip(("rgb", "dist", "weights", "alpha", "d_alpha", "d_rgb_s"), outs[0], outs[1]):
        assert torch.equal(a, b), name
