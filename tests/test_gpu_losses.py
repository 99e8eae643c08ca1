"""GPU parity, stages 4+5 (rgb / depth / reprojection losses, Chamfer) against the reference's golden outputs
and the CPU oracle.  Index outputs are compared bit-exactly."""
import pytest
import torch
import torch.nn.functional as F

from util import assert_close, dev, syn

pytestmark = pytest.mark.gpu


def _loss():
    from model.losses import Loss
    return Loss(syn.default_cfg()["training"])


def test_rgb_loss_vs_reference(golden):
    from model import _ops
    g = golden("losses")["rgb"]
    pred = g["pred"].to(dev()).requires_grad_(True)
    out = _ops.RgbLoss.apply(pred, g["gt"].to(dev()))
    assert_close(out[0], g["l1"], 1e-5, 1e-7, "l1")
    assert_close(out[1], g["l2"], 1e-5, 1e-7, "l2")
    assert_close(out[2], g["mse"], 1e-5, 1e-7, "mse")
    pc = g["pred"].clone().requires_grad_(True)
    d = pc - g["gt"]
    (0.7 * d.abs().sum() / 50 + 0.2 * (d * d).sum() / 50 + 0.1 * F.mse_loss(pc, g["gt"])).backward()
    (0.7 * out[0] + 0.2 * out[1] + 0.1 * out[2]).backward()
    assert_close(pred.grad, pc.grad, 1e-5, 1e-7, "d pred")


@pytest.mark.parametrize("n", [50, 51])
def test_depth_losses_vs_reference(golden, n):
    from model import _ops
    g = golden("losses")["depth_%d" % n]
    p = g["pred"].to(dev()).requires_grad_(True)
    q = g["gt"].to(dev()).requires_grad_(True)
    assert_close(_ops.DepthL1.apply(p, q, None), g["l1"], 1e-5, 1e-7, "l1")
    inv = _ops.DepthInvariant.apply(p, q, None)
    assert_close(inv, g["inv"], 1e-5, 1e-7, "invariant")
    inv.backward()
    assert_close(p.grad, g["inv_grad_pred"], 1e-4, 1e-6, "d pred")
    assert_close(q.grad, g["inv_grad_gt"], 1e-4, 1e-6, "d gt")
    # lower median + index, like torch.median
    val, idx = _ops.median_select(g["pred"].to(dev()))
    tv, ti = torch.median(g["pred"], dim=0)
    assert float(val) == float(tv) and int(idx) == int(ti)


def test_depth_losses_masked_equal_compacted():
    """dense + mask (our asynchronous form) == the reference's boolean-indexed form."""
    from model import _ops
    from oracle import nope_oracle as O
    g = torch.Generator().manual_seed(12)
    p, q = torch.rand(4099, generator=g) * 5, torch.rand(4099, generator=g) * 5
    m = torch.rand(4099, generator=g) > 0.3
    pc, qc = p.clone().requires_grad_(True), q.clone().requires_grad_(True)
    want_l1 = O.depth_l1(pc[m], qc[m])
    want_inv = O.depth_invariant(pc[m], qc[m])
    (want_l1 + want_inv).backward()
    pd, qd = p.to(dev()).requires_grad_(True), q.to(dev()).requires_grad_(True)
    l1 = _ops.DepthL1.apply(pd, qd, m.to(dev()))
    inv = _ops.DepthInvariant.apply(pd, qd, m.to(dev()))
    assert_close(l1, want_l1, 1e-5, 1e-7, "l1")
    assert_close(inv, want_inv, 1e-5, 1e-7, "invariant")
    (l1 + inv).backward()
    assert_close(pd.grad, pc.grad, 1e-4, 1e-7, "d pred")
    assert_close(qd.grad, qc.grad, 1e-4, 1e-7, "d gt")
    assert float(pd.grad[~m.to(dev())].abs().max()) == 0.0


def test_median_select_sizes():
    from model import _ops
    g = torch.Generator().manual_seed(13)
    for n in (1, 2, 3, 1000, 65536, 100001):
        x = torch.randn(n, generator=g)
        val, idx = _ops.median_select(x.to(dev()))
        tv, _ = torch.median(x, dim=0)
        assert float(val) == float(tv), n
        assert float(x[int(idx)]) == float(tv)


def test_photometric_vs_reference(golden):
    from model import _ops
    from oracle import nope_oracle as O
    g = golden("losses")["rgb_s"]
    a = g["rgb1"].to(dev()).requires_grad_(True)
    b = g["rgb2"].to(dev()).requires_grad_(True)
    l = _ops.Photometric.apply(a, b, g["valid"].to(dev()))
    assert_close(l, g["loss"], 1e-5, 1e-7, "loss")
    ac, bc = g["rgb1"].clone().requires_grad_(True), g["rgb2"].clone().requires_grad_(True)
    O.photometric_loss(ac, bc, g["valid"]).backward()
    l.backward()
    assert_close(a.grad, ac.grad, 1e-5, 1e-8, "d rgb1")
    assert_close(b.grad, bc.grad, 1e-5, 1e-8, "d rgb2")
    empty = _ops.Photometric.apply(a, b, torch.zeros_like(g["valid"]).to(dev()))
    assert float(empty) == 0.0  # the reference's 'invalid mask' branch


def test_chamfer_vs_reference(golden):
    L = _loss()
    c = golden("losses")["chamfer"]
    X, Y = c["X"].to(dev()), c["Y"].to(dev())
    assert torch.equal(L.comp_closest_pts_idx_with_split(X, Y).cpu(), c["idx_xy"])
    assert torch.equal(L.comp_closest_pts_idx_with_split(Y, X).cpu(), c["idx_yx"])
    assert torch.equal(L.comp_closest_pts_idx_with_split(c["lat_src"].to(dev()), c["lat_dst"].to(dev())).cpu(), c["lat_idx"])
    Xr = c["X"].t().contiguous().to(dev()).requires_grad_(True)
    Yr = c["Y"].t().contiguous().to(dev()).requires_grad_(True)
    loss = L.get_pc_loss(Xr.unsqueeze(0), Yr.unsqueeze(0))
    assert_close(loss, c["loss"], 1e-5, 1e-7, "chamfer")
    loss.backward()
    assert_close(Xr.grad, c["grad_X"].t(), 1e-4, 1e-7, "d X")
    assert_close(Yr.grad, c["grad_Y"].t(), 1e-4, 1e-7, "d Y")
    assert torch.isfinite(Xr.grad).all()


@pytest.fixture(params=["tcgen05", "cuda_cores", "mixed"])
def chamfer_pass1(request):
    """Screening pass on the tensor cores (production) or on the CUDA cores (the probe it is checked against)."""
    from model import _cabi as K
    old = K.lib().nnb_debug_chamfer_pass1({"tcgen05": 0, "cuda_cores": 1, "mixed": 2}[request.param])
    yield request.param
    K.lib().nnb_debug_chamfer_pass1(old)


@pytest.mark.parametrize("n,m", [(1, 1), (5, 4099), (3000, 2000), (28830, 28830), (60000, 700)])
def test_chamfer_indices_bit_exact_vs_oracle(n, m, chamfer_pass1):
    from model import _ops
    from oracle import nope_oracle as O
    g = torch.Generator().manual_seed(n + m)
    src, dst = torch.rand(n, 3, generator=g) * 2 - 1, torch.rand(m, 3, generator=g) * 2 - 1
    if n >= 3000:
        src = (src * 64).round() / 64   # quantised coordinates: many exact fp32 distance ties
        dst = (dst * 64).round() / 64
    idx, dist = _ops.chamfer_nn(src.to(dev()), dst.to(dev()), want_dist=True)
    sub = slice(0, min(n, 4096))        # the CPU oracle needs 12*S*D bytes: check a slice at the big size
    want = O.nearest_idx(src[sub].t().contiguous(), dst.t().contiguous(), slice_len=256)
    assert torch.equal(idx.cpu()[sub], want)
    wd = torch.linalg.norm(src[sub] - dst[want], dim=1)
    assert_close(dist[sub], wd, 1e-6, 1e-7, 'nn distance')


def test_chamfer_depth_clouds_both_paths_agree(chamfer_pass1):
    """Back-projected depth clouds at training size (two smooth surfaces 28,830 points each, a small rigid offset
    apart -- dense near-ties between neighbouring pixels): every index and distance of the full problem must agree
    with the brute-force oracle on a slice and be identical between the two screening passes."""
    from model import _ops, _cabi as K
    from oracle import nope_oracle as O
    g = torch.Generator().manual_seed(3)
    v, u = torch.meshgrid(torch.linspace(-0.75, 0.75, 155), torch.linspace(-1, 1, 186), indexing="ij")
    z = 2.0 + 0.3 * torch.sin(3 * u) * torch.cos(2 * v)
    a = torch.stack([u * z, v * z, -z], -1).reshape(-1, 3)
    b = a + torch.tensor([0.013, -0.007, 0.02]) + 1e-3 * torch.randn(a.shape, generator=g)
    idx, dist = _ops.chamfer_nn(a.to(dev()), b.to(dev()), want_dist=True)
    sub = slice(11000, 11000 + 4096)
    want = O.nearest_idx(a[sub].t().contiguous(), b.t().contiguous(), slice_len=256)
    assert torch.equal(idx.cpu()[sub], want)
    old = K.lib().nnb_debug_chamfer_pass1(1)
    try:
        idx2, dist2 = _ops.chamfer_nn(a.to(dev()), b.to(dev()), want_dist=True)
    finally:
        K.lib().nnb_debug_chamfer_pass1(old)
    assert torch.equal(idx, idx2) and torch.equal(dist, dist2)


def test_chamfer_source_sharding_matches_one_rank():
    """Data-parallel Chamfer: every rank searches its slice of the source points.  Simulated here with the
    rank contexts of a world of 3 on one GPU: the rank-average of losses and gradients equals the one-rank result."""
    from model import parallel
    L = _loss()
    g = torch.Generator().manual_seed(9)
    X0, Y0 = torch.rand(1, 3001, 3, generator=g).to(dev()), torch.rand(1, 2500, 3, generator=g).to(dev())
    Xf, Yf = X0.clone().requires_grad_(True), Y0.clone().requires_grad_(True)
    full = L.get_pc_loss(Xf, Yf)
    full.backward()
    world = 3
    acc_l, acc_x, acc_y = 0.0, torch.zeros_like(X0), torch.zeros_like(Y0)
    for rank in range(world):
        L.dp = parallel.DataParallelContext(rank, world)
        Xs, Ys = X0.clone().requires_grad_(True), Y0.clone().requires_grad_(True)
        part = L.get_pc_loss(Xs, Ys)
        part.backward()
        acc_l, acc_x, acc_y = acc_l + part.detach() / world, acc_x + Xs.grad / world, acc_y + Ys.grad / world
    assert_close(acc_l, full.detach(), 1e-5, 1e-7, "sharded chamfer loss")
    assert_close(acc_x, Xf.grad, 1e-4, 1e-8, "sharded d X")
    assert_close(acc_y, Yf.grad, 1e-4, 1e-8, "sharded d Y")


def test_chamfer_indices_far_from_origin(chamfer_pass1):
    """Clouds offset far from the origin: the expanded |q|^2 - 2 s.q screen loses ~7 digits to cancellation there;
    the exact re-check must keep the indices bit-exact."""
    from model import _ops
    from oracle import nope_oracle as O
    g = torch.Generator().manual_seed(77)
    off = torch.tensor([30.0, -17.0, 55.0])
    src, dst = torch.rand(2048, 3, generator=g) * 0.5 + off, torch.rand(5000, 3, generator=g) * 0.5 + off
    idx, dist = _ops.chamfer_nn(src.to(dev()), dst.to(dev()), want_dist=True)
    want = O.nearest_idx(src.t().contiguous(), dst.t().contiguous(), slice_len=256)
    assert torch.equal(idx.cpu(), want)


@pytest.mark.parametrize("layout", ["raster", "shuffled", "far_centre"])
def test_chamfer_indices_on_clouds_that_collapse_near_the_camera(layout, chamfer_pass1):
    """Sparse depth maps (the fork's V_KITTI configs) back-project to clouds whose bulk sits in a cap of radius ~0.01 around the
    camera (zero depths clamped to nearest_limit) while one point in six sets the extent: the tensor-core screen's margin carries
    a per-destination share (K slot 15) so that the cap is tested against its own error scale.  Indices must stay bit-exact --
    checked on every source -- for pixel-ordered clouds (stage lists), shuffled ones (plain sweeps), and with the first
    destination (the screen's centre) far from the cap, where the cap points have the LARGEST scaled norms."""
    from model import _ops
    from oracle import nope_oracle as O
    g = torch.Generator().manual_seed(21)
    h, w = 47, 155
    v, u = torch.meshgrid(torch.linspace(-0.3, 0.3, h), torch.linspace(-1, 1, w), indexing="ij")
    dirs = torch.stack([u, v, -torch.ones_like(u)], -1).reshape(-1, 3)

    def cloud(seed, shift):
        gg = torch.Generator().manual_seed(seed)
        far = torch.rand(h * w, generator=gg) < 0.16
        depth = torch.where(far, 0.08 + 0.06 * torch.rand(h * w, generator=gg), torch.full((h * w,), 0.01))
        return dirs * depth.unsqueeze(-1) + shift
    a = cloud(1, torch.tensor([0.012, -0.004, 0.006]))
    b = cloud(2, torch.zeros(3))
    if layout == "shuffled":
        a, b = a[torch.randperm(a.shape[0], generator=g)], b[torch.randperm(b.shape[0], generator=g)]
    if layout == "far_centre":
        b = torch.cat([torch.tensor([[0.9, -0.2, -0.15]]), b])      # dst[0] is the centre of the scaled frame
    from model import _cabi as K
    for src, dst in ((a, b), (b, a)):
        idx, dist = _ops.chamfer_nn(src.to(dev()), dst.to(dev()), want_dist=True)
        old = K.lib().nnb_debug_chamfer_pass1(1)          # every source against the CUDA-core sweeps ...
        try:
            idx_cc, dist_cc = _ops.chamfer_nn(src.to(dev()), dst.to(dev()), want_dist=True)
        finally:
            K.lib().nnb_debug_chamfer_pass1(old)
        assert torch.equal(idx, idx_cc) and torch.equal(dist, dist_cc), "%d of %d indices differ" % (int((idx != idx_cc).sum()), idx.numel())
        sub = torch.arange(0, src.shape[0], 7)            # ... and every seventh against the brute-force oracle on the CPU
        want = O.nearest_idx(src[sub].t().contiguous(), dst.t().contiguous(), slice_len=256)
        assert torch.equal(idx.cpu()[sub], want)
        assert_close(dist[sub.to(dev())], torch.linalg.norm(src[sub] - dst[want], dim=1), 1e-6, 1e-8, "nn distance")


def test_backproject_and_reproject_vs_oracle():
    from model import _ops
    from oracle import nope_oracle as O
    scene = syn.make_scene("tiny", h=37, w=50)
    h, w, h4, w4 = 37, 50, 9, 12
    K = scene["camera_mat"]
    ss = torch.tensor([1.1, 0.02], requires_grad=True)
    depth = scene["depths"][0].clone()
    depth[8, 12] = -5.0  # clamped below nearest_limit
    d = (depth * ss[0] + ss[1]).reshape(1, 1, h, w)
    pc, p_pc = O.backproject_cloud(d, K, (h4, w4), 0.01)
    g = torch.Generator().manual_seed(5)
    r = 0.05 * torch.randn(3, generator=g)
    Rt = O.pose_c2w(r, torch.tensor([0.05, -0.02, 0.3]))
    R_rel, t_rel = Rt[:3, :3].clone().requires_grad_(True), Rt[:3, 3].clone().requires_grad_(True)
    im1 = F.interpolate(scene["imgs"][0:1], (h4, w4), mode="bilinear")
    im2 = F.interpolate(scene["imgs"][1:2], (h4, w4), mode="bilinear")
    rot = pc @ R_rel.t() + t_rel
    behind = (-rot[:, :, 2:] < 0.01).expand_as(rot)
    rot = torch.where(behind, torch.full_like(rot, 0.01), rot)
    xy, valid = O.project(rot, K)
    want1, want2 = O.sample_bilinear(im1, p_pc), O.sample_bilinear(im2, xy)
    probe = torch.randn(h4 * w4, 3, generator=g)
    ((want2[0] * probe).sum() + (pc[0] * probe).sum() * 0.1).backward()

    ssd = ss.detach().to(dev()).requires_grad_(True)
    Rd, td = R_rel.detach().to(dev()).requires_grad_(True), t_rel.detach().to(dev()).requires_grad_(True)
    pc2 = _ops.Backproject.apply(depth.to(dev()), ssd, K.to(dev()), h4, w4, 0.01)
    assert_close(pc2, pc[0], 1e-5, 1e-6, "cloud")
    i1 = _ops.resize_bilinear(scene["imgs"][0].to(dev()), h4, w4)
    i2 = _ops.resize_bilinear(scene["imgs"][1].to(dev()), h4, w4)
    assert_close(i1, im1[0], 1e-5, 1e-6, "resize")
    r1, r2, v = _ops.Reproject.apply(i1, i2, pc2, Rd, td, K.to(dev()), 0.01)
    assert torch.equal(v.cpu(), valid.reshape(-1))
    assert_close(r1, want1[0], 1e-4, 1e-5, "rgb_pc1")
    assert_close(r2, want2[0], 1e-4, 1e-5, "rgb_pc1_proj")
    ((r2 * probe.to(dev())).sum() + (pc2 * probe.to(dev())).sum() * 0.1).backward()
    assert_close(ssd.grad, ss.grad, 1e-3, 1e-5, "d scale/shift")
    assert_close(Rd.grad, R_rel.grad, 1e-3, 1e-5, "d R_rel")
    assert_close(td.grad, t_rel.grad, 1e-3, 1e-5, "d t_rel")


@pytest.mark.parametrize("n1,n2,use_scale", [(1, 1, True), (777, 1030, True), (28830, 28830, True), (500, 333, False)])
def test_align_clouds_vs_reference_expression(n1, n2, use_scale):
    """X = (pc1 / s) @ R^T + t, Y = pc2 / s: the reference's own expression (model/training.py:394-399) evaluated by
    torch in fp64 on the CPU is the checker, forward and all four gradients."""
    from model import _ops
    g = torch.Generator().manual_seed(3 + n1)
    pc1 = torch.randn(n1, 3, generator=g) * 2 + torch.tensor([0.0, 0.0, 4.0])
    pc2 = torch.randn(n2, 3, generator=g) * 2 + torch.tensor([0.0, 0.0, 4.0])
    A = torch.linalg.qr(torch.randn(3, 3, generator=g))[0]
    Rt = torch.eye(4)
    Rt[:3, :3] = A
    Rt[:3, 3] = torch.randn(3, generator=g) * 0.1
    s = torch.tensor(1.3)
    wX, wY = torch.randn(n1, 3, generator=g), torch.randn(n2, 3, generator=g)

    a1, a2, aRt, asc = [t.double().requires_grad_(True) for t in (pc1, pc2, Rt, s)]
    q1, q2 = (a1 / asc, a2 / asc) if use_scale else (a1, a2)
    Xr = q1 @ aRt[:3, :3].transpose(0, 1) + aRt[:3, 3]
    ((Xr * wX.double()).sum() + (q2 * wY.double()).sum()).backward()

    b1, b2, bRt, bsc = [t.to(dev()).requires_grad_(True) for t in (pc1, pc2, Rt, s)]
    X, Y = _ops.AlignClouds.apply(b1, b2, bRt, bsc if use_scale else None)
    ((X * wX.to(dev())).sum() + (Y * wY.to(dev())).sum()).backward()
    assert_close(X, Xr.float(), 1e-5, 1e-6, "X")
    assert_close(Y, q2.float(), 1e-5, 1e-6, "Y")
    assert_close(b1.grad, a1.grad.float(), 1e-5, 1e-6, "d pc1")
    assert_close(b2.grad, a2.grad.float(), 1e-5, 1e-6, "d pc2")
    tol = 2e-5 * float(aRt.grad.abs().max()) + 1e-6      # fp32 sums over the points
    assert_close(bRt.grad, aRt.grad.float(), 1e-4, tol, "d Rt")
    assert float(bRt.grad[3].abs().max()) == 0.0
    if use_scale:
        assert_close(bsc.grad, asc.grad.float(), 1e-4, 2e-5 * float(asc.grad.abs()) + 1e-6, "d scale")
    else:
        assert bsc.grad is None
