"""Edge cases of the hot path: tiny and odd ray counts (the tensor-core kernels work on ray PAIRS per CTA and CTA pairs per
cluster, so 1, 2, 3 rays and counts around a multiple of 128 exercise the dummy tiles), a batch whose depth is invalid everywhere
(model/rendering.py:69-80 masks zero depths; model/losses.py:60-63 then divides by a zero count in the reference -- the kernels
return 0 and a zero gradient instead of NaN), one-point clouds, and an empty data-parallel shard (refused with a clear error)."""
import pytest
import torch

from util import assert_close, build_models, dev, max_err, syn

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("R", [1, 2, 3, 127, 129])
@pytest.mark.parametrize("precision", ["fp32", "bf16"])
def test_render_forward_backward_at_tiny_and_odd_ray_counts(R, precision):
    from model import _cabi, _ops
    from oracle import nope_oracle as O
    g = torch.Generator().manual_seed(100 + R)
    sd = syn.init_mlp_state(42)
    dirs = torch.nn.functional.normalize(torch.randn(R, 3, generator=g) + torch.tensor([0., 0., -2.]), dim=-1)
    origin = torch.tensor([0.1, -0.2, 0.3])
    noise = torch.rand(R, 128, generator=g)
    p_rgb, p_dist = torch.randn(R, 3, generator=g), torch.randn(R, generator=g)
    # oracle (fp32, CPU)
    P = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    z = O.stratified_z(R, 128, 0.01, 10.0, noise)
    pts = origin.view(1, 1, 3) + dirs.view(R, 1, 3) * z.view(R, 128, 1)
    rgb_s, alpha = O.mlp_forward(P, pts.reshape(-1, 3), (-dirs).view(R, 1, 3).expand(R, 128, 3).reshape(-1, 3))
    rgb, dist, w, _ = O.composite(alpha.reshape(R, 128), rgb_s.reshape(R, 128, 3), z)
    ((rgb * p_rgb).sum() + (dist * p_dist).sum()).backward()
    # kernels
    prec = _cabi.PREC_BF16 if precision == "bf16" else _cabi.PREC_FP32
    params = [sd[k].to(dev()).requires_grad_(True) for k in _cabi.PARAM_ORDER]
    packed = _ops.pack_weights([p.detach() for p in params]) if precision == "bf16" else None
    rc = _ops.make_cfg(R, 128, 0.01, 10.0, 1, 0, 0, 1, prec)
    od, dd = origin.to(dev()).requires_grad_(True), dirs.to(dev()).requires_grad_(True)
    rgb2, dist2, w2, a2, z2 = _ops.RenderRays.apply(rc, packed, od, dd, noise.to(dev()), *params)
    ((rgb2 * p_rgb.to(dev())).sum() + (dist2 * p_dist.to(dev())).sum()).backward()
    torch.cuda.synchronize()
    assert torch.equal(z2.cpu(), z), "sample depths must be bit-exact"
    # ... and equal to what the reference-named stand-alone routine returns for the same draws (model/rendering.py:183-198)
    import model as mdl
    rcfg = {"depth_range": [0.01, 10.0], "n_max_network_queries": 64000, "white_background": False, "num_points": 128, "outside_steps": 0}
    z_ref, pts_ref, view_ref = mdl.Renderer(torch.nn.Identity(), rcfg, device=torch.device("cpu"), precision="fp32").sample_uniform(
        origin.view(1, 1, 3).expand(1, R, 3), dirs.view(1, R, 3), torch.linspace(0., 1., 128).view(1, 1, 128).repeat(1, R, 1), True, [0.01, 10.0],
        noise=noise.view(1, R, 128))
    assert torch.equal(z_ref.view(R, 128), z2.cpu()) and torch.equal(pts_ref, pts.reshape(-1, 3)) and torch.equal(view_ref, (-dirs).view(R, 1, 3).expand(R, 128, 3).reshape(-1, 3))
    ftol = 1e-5 if precision == "fp32" else 1e-3
    assert max_err(rgb2, rgb) <= ftol and max_err(dist2, dist) <= 10 * ftol and max_err(w2, w) <= ftol
    gtol = 2e-4 if precision == "fp32" else 0.3      # a handful of rays: single ReLU-mask flips dominate the bf16 gradient
    worst = 0.0
    for k, p in zip(_cabi.PARAM_ORDER, params):
        ref = P[k].grad
        worst = max(worst, float((p.grad.cpu() - ref).norm() / ref.norm().clamp_min(1e-20)))
    print("R=%d %s: rgb %.2e dist %.2e, worst gradient rel-L2 %.2e" % (R, precision, max_err(rgb2, rgb), max_err(dist2, dist), worst))
    assert worst <= gtol
    assert bool(torch.isfinite(dd.grad).all()) and bool(torch.isfinite(od.grad).all())


@pytest.mark.parametrize("depth_loss_type", ["l1", "invariant"])
def test_step_with_no_valid_depth_anywhere(depth_loss_type):
    import model as mdl
    cfg = syn.default_cfg()
    scene = syn.make_scene("tiny")
    n, R = scene["n"], 64
    nerf, renderer, net, pose, dist = build_models(cfg, n, scene, precision="fp32")      # scale 1, shift 0: zero depth stays zero
    tcfg = dict(cfg["training"], n_training_points=R, depth_loss_type=depth_loss_type)
    tr = mdl.Trainer(net, None, tcfg, device=dev(), pose_param_net=pose, distortion_net=dist)
    tr.fixed_ray_idx, tr.fixed_noise = syn.ray_choice(scene["h"], scene["w"], R), syn.jitter(R)
    batch = syn.make_batch(scene, 1, 2)
    batch["img.dpt"] = torch.zeros_like(batch["img.dpt"])
    ld = tr.compute_loss(batch, it=1, epoch=0, scheduling_start=cfg["training"]["scheduling_start"])
    ld["loss"].backward()
    torch.cuda.synchronize()
    assert float(ld["loss_depth"]) == 0.0, "no valid ray: the depth term is 0, not NaN"
    for k in ("loss", "loss_rgb", "loss_pc", "loss_rgb_s"):
        assert float(ld[k]) == float(ld[k]), k
    for p in list(nerf.parameters()) + [pose.r, pose.t]:
        assert p.grad is not None and bool(torch.isfinite(p.grad).all())


def test_single_point_clouds_and_identical_clouds():
    from model import _ops
    import model as mdl
    L = mdl.Loss(dict(syn.default_cfg()["training"]))
    a = torch.tensor([[[0.1, 0.2, 0.3]]], device=dev(), requires_grad=True)
    b = torch.tensor([[[0.4, 0.2, -0.1]]], device=dev(), requires_grad=True)
    loss = L.get_pc_loss(a, b)
    loss.backward()
    assert_close(loss, torch.tensor(2 * 0.5), 1e-6, 1e-7, "two one-point clouds 0.5 apart")
    g = torch.Generator().manual_seed(1)
    X = torch.rand(1, 500, 3, generator=g).to(dev()).requires_grad_(True)
    Y = X.detach().clone().requires_grad_(True)
    same = L.get_pc_loss(X, Y)          # every point coincides with its neighbour: zero loss, ZERO subgradient (torch's norm at 0)
    same.backward()
    assert float(same) == 0.0 and float(X.grad.abs().max()) == 0.0 and float(Y.grad.abs().max()) == 0.0
    idx = _ops.chamfer_nn(X[0].detach(), Y[0].detach())
    assert torch.equal(idx.cpu(), torch.arange(500))


def test_empty_data_parallel_shard_is_refused():
    import model as mdl
    from model import parallel

    class DP(parallel.DataParallelContext):
        pass
    cfg = syn.default_cfg()
    scene = syn.make_scene("tiny")
    nerf, renderer, net, pose, dist = build_models(cfg, scene["n"], scene, precision="fp32")
    tr = mdl.Trainer(net, None, dict(cfg["training"], n_training_points=2), device=dev(), pose_param_net=pose, distortion_net=dist,
                     dp=DP(rank=3, world_size=4))
    with pytest.raises(ValueError, match="empty shard"):
        tr.compute_loss(syn.make_batch(scene, 1, 2), it=1, epoch=0, scheduling_start=cfg["training"]["scheduling_start"])
