"""GPU parity, stage 1 (pose, 4x4, distortion, pixel grid, ray selection, ray generation) against the golden
fixtures produced by the reference and against the CPU oracle.  Calls go through the C ABI (ctypes)."""
import pytest
import torch

from util import assert_close, dev, syn, to_dev

pytestmark = pytest.mark.gpu


def test_pose_forward_backward_vs_reference(golden):
    from model import _ops
    g = golden("stages")["pose"]
    n = g["r"].shape[0]
    for i in range(n):
        r = g["r"].clone().to(dev()).requires_grad_(True)
        t = g["t"].clone().to(dev()).requires_grad_(True)
        c = _ops.PoseC2W.apply(r, t, None, i)
        assert_close(c, g["c2w"][i], 1e-5, 1e-6, "c2w[%d]" % i)
        m = _ops.PoseC2W.apply(r, t, g["init"].to(dev()), i)
        assert_close(m, g["c2w_init"][i], 1e-5, 2e-6, "c2w_init[%d]" % i)
        (m * g["probe"].to(dev())).sum().backward()
        # the reference's fp32 autograd loses ~1e-5 relative in (th cos th - sin th) at small th; the kernel is fp64 inside
        assert_close(r.grad[i], g["grad_r"][i], 1e-4, 2e-5, "grad_r[%d]" % i)
        assert_close(t.grad[i], g["grad_t"][i], 1e-5, 1e-6, "grad_t[%d]" % i)
        other = [j for j in range(n) if j != i]
        assert float(r.grad[other].abs().max()) == 0.0
    # Exp(0) is exactly the identity with a finite gradient
    z = torch.zeros(1, 3, device=dev(), requires_grad=True)
    c = _ops.PoseC2W.apply(z, torch.zeros(1, 3, device=dev()), None, 0)
    assert torch.equal(c.cpu(), torch.eye(4))
    c[:3, :3].sum().backward()
    assert torch.isfinite(z.grad).all()


def test_learnpose_module_matches_oracle():
    import model as mdl
    from oracle import nope_oracle as O
    scene = syn.make_scene("tiny")
    r0, t0 = syn.init_pose_params(scene["n"])
    pose = mdl.LearnPose(scene["n"], True, True, {}, init_c2w=scene["poses_gt"].clone()).to(dev())
    with torch.no_grad():
        pose.r.copy_(r0)
        pose.t.copy_(t0)
    for i in range(scene["n"]):
        assert_close(pose(torch.tensor([i])), O.pose_c2w(r0[i], t0[i], scene["poses_gt"][i]), 1e-5, 1e-6, "pose %d" % i)
    with pytest.raises(IndexError):
        pose(scene["n"])


def test_mat4_inverse_and_grad():
    from model import _ops
    g = torch.Generator().manual_seed(3)
    for _ in range(4):
        m = torch.eye(4) + 0.3 * torch.randn(4, 4, generator=g)
        probe = torch.randn(4, 4, generator=g)
        mc = m.clone().requires_grad_(True)
        (torch.inverse(mc) * probe).sum().backward()
        md = m.to(dev()).requires_grad_(True)
        inv = _ops.Mat4Inverse.apply(md)
        assert_close(inv, torch.inverse(m), 1e-5, 1e-6, "inverse")
        (inv * probe.to(dev())).sum().backward()
        assert_close(md.grad, mc.grad, 1e-4, 1e-5, "inverse grad")


def test_distortion_vs_reference(golden):
    import model as mdl
    g = golden("stages")["distortion"]
    net = mdl.Learn_Distortion(4, True, True, syn.default_cfg()).to(dev())
    with torch.no_grad():
        net.global_scales.copy_(g["scales"])
        net.global_shifts.copy_(g["shifts"])
    for i in range(4):
        s, sh = net(i)
        assert abs(float(s) - g["out"][i][0]) < 1e-7 and abs(float(sh) - g["out"][i][1]) < 1e-7
    # the scale gradient passes for camera 0, is a zero row for the clamped camera (1) and ABSENT for the pinned one (3): there the
    # reference's scale is a constant, `global_scales.grad` stays None and torch.optim.Adam skips the tensor (model/distortions.py:
    # 21-24; tests/test_gpu_trajectory.py shows what a zero gradient instead of none does to the trajectory)
    for i, want in ((0, 1.0), (1, 0.0), (3, None)):
        net.zero_grad(set_to_none=True)
        s, sh = net(i)
        (s.sum() + 2 * sh.sum()).backward()
        if want is None:
            assert net.global_scales.grad is None
        else:
            assert float(net.global_scales.grad[i]) == want
        assert float(net.global_shifts.grad[i]) == 2.0
    # nothing downstream asked for a gradient: no gradient, not a zero one
    net.zero_grad(set_to_none=True)
    ss = net.scale_shift(0)
    (ss.detach().sum() + net.global_shifts[2].sum()).backward()
    assert net.global_scales.grad is None


def test_pixel_grid_bit_exact(golden):
    from model import _ops
    g = golden("stages")["pixel_grid"]
    loc, sc = _ops.pixel_grid(g["h"], g["w"], dev())
    assert torch.equal(loc.cpu().unsqueeze(0), g["loc"])
    assert torch.equal(sc.cpu().unsqueeze(0), g["scaled"])
    from oracle import nope_oracle as O
    for h, w in ((375, 1242), (93, 310), (120, 160)):
        _, sc = _ops.pixel_grid(h, w, dev())
        assert torch.equal(sc.cpu().unsqueeze(0), O.pixel_grid(h, w)[1])


def test_gather_rays_bit_exact_and_grad():
    from model import _ops
    from oracle import nope_oracle as O
    scene = syn.make_scene("tiny")
    h, w = scene["h"], scene["w"]
    idx = syn.ray_choice(h, w, 64)
    ss = torch.tensor([1.05, 0.01], requires_grad=True)
    depth_full = scene["depths"][1] * ss[0] + ss[1]
    want_d = depth_full.reshape(-1)[idx]
    want_rgb = scene["imgs"][1].reshape(3, -1).t()[idx]
    want_p = O.pixel_grid(h, w)[1][0, idx]
    ssd = ss.detach().to(dev()).requires_grad_(True)
    p, rgb, d = _ops.GatherRays.apply(idx.to(dev()), scene["imgs"][1].to(dev()), scene["depths"][1].to(dev()), ssd, h, w)
    assert torch.equal(p.cpu(), want_p) and torch.equal(rgb.cpu(), want_rgb) and torch.equal(d.cpu(), want_d.detach())
    probe = torch.randn(64, generator=torch.Generator().manual_seed(1))
    (want_d * probe).sum().backward()
    (d * probe.to(dev())).sum().backward()
    assert_close(ssd.grad, ss.grad, 1e-5, 1e-6, "d scale/shift")


@pytest.mark.parametrize("normalise", [True, False])
def test_raygen_vs_oracle(normalise):
    from model import _ops
    from oracle import nope_oracle as O
    scene = syn.make_scene("tiny")
    h, w = scene["h"], scene["w"]
    idx = syn.ray_choice(h, w, 64)
    pixels = O.pixel_grid(h, w)[1][:, idx]
    depth = scene["depths"][1].reshape(1, -1, 1)[:, idx].clone()
    depth[0, 3, 0] = 0.0
    r0, t0 = syn.init_pose_params(scene["n"])
    c2w = O.pose_c2w(r0[1], t0[1], scene["poses_gt"][1])
    Wc = torch.inverse(c2w).unsqueeze(0).requires_grad_(True)
    dc = depth.clone().requires_grad_(True)
    o, d, nrm, dgt, mask = O.ray_setup(pixels, dc, scene["camera_mat"], Wc, scene["scale_mat"], normalise_ray=normalise)
    Wd = Wc.detach().to(dev()).requires_grad_(True)
    dd = depth.to(dev()).requires_grad_(True)
    Kd = scene["camera_mat"].to(dev()).requires_grad_(True)
    o2, d2, nrm2, dgt2, mask2 = _ops.RayGen.apply(pixels.to(dev()), dd, Kd, Wd, scene["scale_mat"].to(dev()), normalise)
    assert_close(o2, o, 1e-5, 1e-6, "origin")
    assert_close(d2, d, 1e-5, 1e-6, "dirs")
    assert_close(nrm2, nrm, 1e-5, 1e-6, "norm")
    assert torch.equal(mask2.cpu(), mask)
    assert_close(dgt2[mask2], dgt[mask], 1e-5, 1e-6, "d_gt")
    g = torch.Generator().manual_seed(2)
    p_o, p_d, p_g = torch.randn(3, generator=g), torch.randn(64, 3, generator=g), torch.randn(64, generator=g) * mask
    ((o * p_o).sum() + (d * p_d).sum() + (dgt[mask] * p_g[mask]).sum()).backward()
    ((o2 * p_o.to(dev())).sum() + (d2 * p_d.to(dev())).sum() + (dgt2[mask2] * p_g.to(dev())[mask2]).sum()).backward()
    assert_close(Wd.grad, Wc.grad, 1e-4, 1e-4, "d world_mat")
    fin = torch.isfinite(depth.reshape(-1))
    assert_close(dd.grad.reshape(-1)[fin], dc.grad.reshape(-1)[fin], 1e-4, 1e-5, "d depth")
    assert torch.isfinite(Kd.grad).all()
