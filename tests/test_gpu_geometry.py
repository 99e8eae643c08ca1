"""GPU parity of the geometry path: OfficialStaticNerf.gradient / infer_occ, Renderer.ray_marching
(+ secant) and Renderer.phong_renderer against fixtures made by the reference (tests/golden/geometry.pt,
oracle/make_golden.py::geometry_vectors).  fp32 kernels; tolerances: field values 1e-4, input gradient 1e-3 relative to
its scale, surface distances 1e-3 (the secant iteration divides by small occupancy differences), shaded colours 5e-3."""
import pytest
import torch

from util import assert_close, dev, max_err, syn

pytestmark = pytest.mark.gpu


def _geometry_state():
    sd = {k: v.clone() for k, v in syn.init_mlp_state(42).items()}
    sd["fc_density.weight"] = sd["fc_density.weight"] * 200.0     # oracle/make_golden.py::geometry_state
    sd["fc_density.bias"] = torch.tensor([0.86])
    return sd


def _renderer():
    import model as mdl
    cfg = syn.default_cfg()
    nerf = mdl.OfficialStaticNerf(cfg)
    nerf.load_state_dict(_geometry_state())
    return cfg, nerf, mdl.Renderer(nerf, dict(cfg["rendering"], precision="fp32"), device=dev())


def test_density_gradient_and_occupancy_vs_reference(golden):
    g = golden("geometry")["gradient"]
    _, nerf, _ = _renderer()
    pts = g["pts"].to(dev())
    grad = nerf.gradient(pts, 0)
    assert grad.shape == g["grad"].shape
    scale = float(g["grad"].abs().max())
    print("density gradient max err %.3e (scale %.3e)" % (max_err(grad, g["grad"]), scale))
    assert max_err(grad, g["grad"]) <= 1e-3 * scale
    _, raw = nerf.infer_occ(pts)
    assert_close(raw, g["raw"], 1e-4, 1e-4, "raw density")
    occ = nerf(pts, only_occupancy=True)
    assert_close(occ, g["occ"], 1e-4, 2e-5, "occupancy")
    # more points than one chunk of the kernel: chunking must not change values
    big = pts.repeat(500, 1)
    gb = nerf.gradient(big, 0)
    assert torch.equal(gb[:pts.shape[0]], gb[-pts.shape[0]:]) and max_err(gb[:pts.shape[0]], grad) == 0.0


def test_ray_marching_and_phong_vs_reference(golden):
    g = golden("geometry")["phong"]
    cfg, nerf, renderer = _renderer()
    scene = syn.make_scene("tiny")
    cam = g["camera_world"].to(dev())
    dirs = g["ray_dirs"].to(dev())
    R = dirs.shape[0]
    nerf.eval()
    d = renderer.ray_marching(cam.view(1, 1, 3).expand(1, R, 3), dirs.unsqueeze(0), nerf, n_secant_steps=8, n_steps=[512, 513],
                              rad=cfg["rendering"]["radius"])
    ref = g["d_i"]
    same_class = (torch.isfinite(d.cpu()) == torch.isfinite(ref)) & ((d.cpu() == 0) == (ref == 0))
    assert same_class.float().mean() >= 0.98, "hit / miss / occupied classification"
    ok = same_class & torch.isfinite(ref)
    err = (d.cpu()[ok] - ref[ok]).abs().max()
    print("surface distance max err %.3e over %d rays" % (float(err), int(ok.sum())))
    assert float(err) <= 1e-3
    out = renderer(g["pixels"].to(dev()), None, scene["camera_mat"].to(dev()), g["world_mat"].to(dev()), scene["scale_mat"].to(dev()),
                   "phong_renderer", it=0)
    assert set(out) == {"rgb", "normal", "rgb_surf"} and out["normal"] is None
    m = ok[0]
    print("phong rgb max err %.3e, surface colour max err %.3e" % (max_err(out["rgb"][0][m], g["rgb"][0][m]),
                                                                  max_err(out["rgb_surf"][0][m], g["rgb_surf"][0][m])))
    assert max_err(out["rgb"][0][m], g["rgb"][0][m]) <= 5e-3
    assert max_err(out["rgb_surf"][0][m], g["rgb_surf"][0][m]) <= 5e-3


def test_sphere_intersection_and_misses():
    """Rays that miss the sphere or start inside occupied space (model/rendering.py:376-385 conventions: inf / 0)."""
    from model import geometry
    from oracle import nope_oracle as O
    g = torch.Generator().manual_seed(4)
    cam = torch.tensor([0.3, -0.2, 6.0])
    dirs = torch.nn.functional.normalize(torch.randn(200, 3, generator=g) + torch.tensor([0., 0., -1.5]), dim=-1)
    want, wmask = O.sphere_intersection(cam, dirs, 4.0)
    got, gmask = geometry.get_sphere_intersection(cam.to(dev()), dirs.to(dev()), 4.0)
    assert torch.equal(gmask.cpu(), wmask) and 0 < int(wmask.sum()) < 200
    assert_close(got, want, 1e-5, 1e-5, "sphere intersections")


def test_normal_loss_output_vs_oracle():
    """rendering.normal_loss (model/rendering.py:127-135): out_dict['normal'] = |n(p) - n(p + jitter)| at the depth-map surface
    points, normals from the density gradient.  The jitter is the first device draw after the (injected) stratified noise."""
    import copy
    import model as mdl
    from oracle import nope_oracle as O
    cfg = copy.deepcopy(syn.default_cfg())
    cfg["rendering"]["normal_loss"] = True
    scene = syn.make_scene("tiny")
    sd = _geometry_state()
    nerf = mdl.OfficialStaticNerf(cfg)
    nerf.load_state_dict(sd)
    renderer = mdl.Renderer(nerf, dict(cfg["rendering"], precision="fp32"), device=dev())
    R = 40
    idx = syn.ray_choice(scene["h"], scene["w"], R)
    pixels = O.pixel_grid(scene["h"], scene["w"])[1][:, idx]
    depth = scene["depths"][1].reshape(1, -1, 1)[:, idx].clone()
    depth[0, 5, 0] = 0.0                                        # one masked ray
    world = torch.inverse(scene["poses_gt"][1]).unsqueeze(0)
    noise = syn.jitter(R)
    torch.manual_seed(77)
    out = renderer(pixels.to(dev()), depth.to(dev()), scene["camera_mat"].to(dev()), world.to(dev()), scene["scale_mat"].to(dev()),
                   "nope_nerf", add_noise=True, eval_=False, noise=noise.to(dev()))
    assert out["normal"] is not None and out["normal"].shape == (R - 1,)
    torch.manual_seed(77)
    jit = torch.rand(R - 1, 3, device=dev()).cpu()              # the same draw as rand_like(surface_points)
    o, d, _, d_gt, mask = O.ray_setup(pixels, depth, scene["camera_mat"], world, scene["scale_mat"])
    surf = (o.reshape(1, 3) + d * d_gt.reshape(-1, 1))[mask]
    pp = torch.cat([surf, surf + (jit - 0.5) * 0.01], 0)
    g = O.density_gradient(sd, pp)[:, 0, :]
    n = g / (g.norm(2, dim=1).unsqueeze(-1) + 1e-5)
    want = torch.norm(n[:R - 1] - n[R - 1:], dim=-1)
    print("normal-loss output max err %.3e (values up to %.3e)" % (max_err(out["normal"], want), float(want.max())))
    assert_close(out["normal"], want, 2e-3, 2e-4, "diff_norm")


def test_fused_proposals_bf16_vs_fp32_marching():
    """bf16 renderer: the 512 proposals per ray go through the fused tensor-core kernel (geometry._proposals_fused), the secant
    steps stay fp32.  Against the fp32 point-kernel marcher: same hit / miss / occupied class on >= 97% of the rays (a bf16
    occupancy within ~1e-2 of tau at a proposal can move the bracket by one step), and on those the refined distance agrees to
    2e-3 (the secant iteration converges on the fp32 field from either bracket)."""
    import model as mdl
    from model import geometry
    cfg, nerf, r32 = _renderer()
    r16 = mdl.Renderer(nerf, dict(cfg["rendering"], precision="bf16"), device=dev())
    nerf.eval()
    g = torch.Generator().manual_seed(5)
    R = 3000
    cam = torch.tensor([0.1, -0.2, -0.9], device=dev())
    dirs = torch.nn.functional.normalize(torch.randn(R, 3, generator=g) * torch.tensor([0.4, 0.4, 0.1]) + torch.tensor([0., 0., 1.]), dim=-1).to(dev())
    rad = cfg["rendering"]["radius"]
    with torch.no_grad():
        d32 = geometry.ray_marching(nerf, cam, dirs, rad=rad)
        d16 = geometry.ray_marching(nerf, cam, dirs, rad=rad, renderer=r16)
        # the proposal values themselves: fused bf16 vs fp32 point kernels
        inter, _ = geometry.get_sphere_intersection(cam, dirs, r=rad)
        u = torch.linspace(0, 1, steps=512, device=dev()).view(1, 512)
        d_prop = inter[:, 1:2] * u
        v16 = geometry._proposals_fused(r16, cam, dirs, d_prop, 0.5)
        p = cam.view(1, 1, 3) + dirs.unsqueeze(1) * d_prop.unsqueeze(-1)
        v32 = geometry._occupancy(nerf, p.reshape(-1, 3), 0.5, 3500000).view(R, 512)
    print("proposal occupancy bf16 vs fp32: max err %.3e" % max_err(v16, v32))
    assert max_err(v16, v32) <= 6e-2
    same = (torch.isfinite(d16) == torch.isfinite(d32)) & ((d16 == 0) == (d32 == 0))
    hit = same & torch.isfinite(d32) & (d32 > 0)
    print("class agreement %.4f, hits %d, distance max err %.3e, median %.3e" % (
        float(same.float().mean()), int(hit.sum()), float((d16[hit] - d32[hit]).abs().max()) if hit.any() else 0.0,
        float((d16[hit] - d32[hit]).abs().median()) if hit.any() else 0.0))
    assert float(same.float().mean()) >= 0.97
    assert hit.sum() > 100
    close = (d16[hit] - d32[hit]).abs() <= 2e-3
    assert float(close.float().mean()) >= 0.97
