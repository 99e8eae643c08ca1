"""GPU parity of the tcgen05 (bf16 operand, fp32 accumulate) render path against the CPU oracle.
Tolerance (north_star: "<= 1e-2 abs rgb if bf16 tensor cores are used"): rgb 1e-2 abs, alpha/weights 1e-2 abs,
expected distance 1e-2 relative to the far plane; sample depths stay bit-exact."""
import pytest
import torch

from util import dev, max_err, syn

pytestmark = pytest.mark.gpu


def _oracle_forward(R, seed=0, noise_on=True):
    from oracle import nope_oracle as O
    g = torch.Generator().manual_seed(100 + seed)
    o = torch.randn(3, generator=g) * 0.3
    d = torch.nn.functional.normalize(torch.randn(R, 3, generator=g) + torch.tensor([0., 0., -2.]), dim=-1)
    noise = torch.rand(R, 128, generator=g) if noise_on else None
    sd = syn.init_mlp_state(42)
    with torch.no_grad():
        z = O.stratified_z(R, 128, 0.01, 10.0, noise)
        pts = (o.reshape(1, 1, 3) + d.unsqueeze(1) * z.unsqueeze(-1)).reshape(-1, 3)
        view = (-d).unsqueeze(1).expand(R, 128, 3).reshape(-1, 3)
        rgb_s, alpha = O.mlp_forward(sd, pts, view)
        rgb, dist, w, _ = O.composite(alpha.reshape(R, 128), rgb_s.reshape(R, 128, 3), z)
    return sd, o, d, noise, z, rgb, dist, w, alpha.reshape(R, 128)


@pytest.mark.parametrize("R,noise_on", [(64, True), (7, False), (601, True)])
def test_tc_forward_vs_oracle(R, noise_on):
    from model import _cabi, _ops
    sd, o, d, noise, z, rgb, dist, w, alpha = _oracle_forward(R, seed=R, noise_on=noise_on)
    params = [sd[k].to(dev()) for k in _cabi.PARAM_ORDER]
    packed = _ops.pack_weights(params)
    rc = _ops.make_cfg(R, 128, 0.01, 10.0, 1, 0, 0, 1, _cabi.PREC_BF16)
    rgb2, dist2, w2, a2, z2 = _ops.render_nograd(rc, params, packed, o.to(dev()), d.to(dev()),
                                                 None if noise is None else noise.to(dev()))
    torch.cuda.synchronize()
    assert torch.equal(z2.cpu(), z), "sample depths must be bit-exact"
    errs = dict(rgb=max_err(rgb2, rgb), dist=max_err(dist2, dist), alpha=max_err(a2, alpha), weights=max_err(w2, w))
    print("tc forward R=%d max abs err: %s" % (R, {k: "%.3e" % v for k, v in errs.items()}))
    assert errs["rgb"] <= 1e-2 and errs["alpha"] <= 1e-2 and errs["weights"] <= 1e-2
    assert errs["dist"] <= 1e-2 * 10.0
    # no-aux call (render loops) gives the same rgb/dist
    rgb3, dist3, _, _, _ = _ops.render_nograd(rc, params, packed, o.to(dev()), d.to(dev()),
                                              None if noise is None else noise.to(dev()), want_aux=False)
    assert torch.equal(rgb3, rgb2) and torch.equal(dist3, dist2)


def _reference_grads(mlp, sd, o0, d0, noise, p_rgb, p_dist, R):
    from oracle import nope_oracle as O
    o, d = o0.clone().requires_grad_(True), d0.clone().requires_grad_(True)
    P = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    z = O.stratified_z(R, 128, 0.01, 10.0, noise)
    pts = (o.reshape(1, 1, 3) + d.unsqueeze(1) * z.unsqueeze(-1)).reshape(-1, 3)
    view = (-d).unsqueeze(1).expand(R, 128, 3).reshape(-1, 3)
    rgb_s, alpha = mlp(P, pts, view)
    rgb, dist, w, _ = O.composite(alpha.reshape(R, 128), rgb_s.reshape(R, 128, 3), z)
    ((rgb * p_rgb).sum() + (dist * p_dist).sum()).backward()
    return rgb.detach(), dist.detach(), {k: P[k].grad for k in P}, o.grad, d.grad


def _rel_l2(a, b):
    return float((a.detach().cpu() - b).norm() / b.norm().clamp_min(1e-20))


@pytest.mark.parametrize("R", [64, 33])
def test_tc_backward_vs_oracle(R):
    """Gradients of the bf16 tensor-core path.
    (1) Tight: against the bf16-emulating oracle (oracle/bf16_emulation.py: same rounding points, fp32 accumulate):
        relative L2 error <= 5e-2 per gradient tensor (measured 4e-3 .. 2.5e-2; the rest is a handful of mask flips from summation order) -- this is the kernel-correctness check.
    (2) Loose: against the fp32 oracle: relative L2 <= 0.18 at these 33 / 64 rays (0.06 at 4,096 rays, tests/test_gpu_cfg2.py) -- the documented bf16-vs-fp32 deviation, dominated by
        ReLU masks flipping for pre-activations within bf16 rounding distance of zero."""
    from model import _cabi, _ops
    from oracle import nope_oracle as O
    from oracle.bf16_emulation import mlp_forward_bf16emu
    g = torch.Generator().manual_seed(200 + R)
    o = torch.randn(3, generator=g) * 0.3
    d = torch.nn.functional.normalize(torch.randn(R, 3, generator=g) + torch.tensor([0., 0., -2.]), dim=-1)
    noise = torch.rand(R, 128, generator=g)
    sd = syn.init_mlp_state(42)
    p_rgb, p_dist = torch.randn(R, 3, generator=g) / R, torch.randn(R, generator=g) / R
    ref32 = _reference_grads(O.mlp_forward, sd, o, d, noise, p_rgb, p_dist, R)
    refbf = _reference_grads(mlp_forward_bf16emu, sd, o, d, noise, p_rgb, p_dist, R)

    rc = _ops.make_cfg(R, 128, 0.01, 10.0, 1, 0, 0, 1, _cabi.PREC_BF16)
    params = [sd[k].to(dev()).requires_grad_(True) for k in _cabi.PARAM_ORDER]
    packed = _ops.pack_weights([p.detach() for p in params])
    od, dd = o.to(dev()).requires_grad_(True), d.to(dev()).requires_grad_(True)
    rgb2, dist2, w2, a2, z2 = _ops.RenderRays.apply(rc, packed, od, dd, noise.to(dev()), *params)
    ((rgb2 * p_rgb.to(dev())).sum() + (dist2 * p_dist.to(dev())).sum()).backward()
    torch.cuda.synchronize()
    assert max_err(rgb2, ref32[0]) <= 1e-2
    print("tc R=%d forward vs bf16-emulation: rgb %.2e dist %.2e" % (R, max_err(rgb2, refbf[0]), max_err(dist2, refbf[1])))
    # fp32 oracle: measured 0.090 (R = 64) and 0.128 (R = 33) on B200 -- a handful of rays, so single ReLU-mask flips weigh heavily;
    # at 4,096 rays the same comparison gives 0.029 (tests/test_gpu_cfg2.py asserts 0.06)
    for name, ref, tol in (("bf16-emulation", refbf, 5e-2), ("fp32 oracle", ref32, 0.18)):
        errs = {k: _rel_l2(p.grad, ref[2][k]) for k, p in zip(_cabi.PARAM_ORDER, params)}
        errs["d_origin"], errs["d_dirs"] = _rel_l2(od.grad, ref[3]), _rel_l2(dd.grad, ref[4])
        print("tc backward R=%d vs %s: worst rel-L2 %.3e" % (R, name, max(errs.values())))
        print("   ", {k: "%.1e" % v for k, v in errs.items()})
        assert max(errs.values()) <= tol, name


def _full_size_inputs(R, seed=5):
    from model import _cabi
    g = torch.Generator().manual_seed(seed)
    sd = syn.init_mlp_state(42)
    params = [sd[k].to(dev()).requires_grad_(True) for k in _cabi.PARAM_ORDER]
    o = (torch.randn(3, generator=g) * 0.3).to(dev()).requires_grad_(True)
    d = torch.nn.functional.normalize(torch.randn(R, 3, generator=g) + torch.tensor([0., 0., -2.]), dim=-1).to(dev()).requires_grad_(True)
    noise = torch.rand(R, 128, generator=g).to(dev())
    return params, o, d, noise


def test_tc_full_size_properties_cfg2():
    """BASELINE.json cfg2 size (4,096 rays x 128 samples), where the CPU oracle takes minutes: size-independent properties.
    (1) ranges: rgb in [0,1], weights >= 0 with sum <= 1, near <= expected distance / sum(w) <= far;
    (2) rays are independent: rendering a permutation of the rays gives the permuted result bit for bit, in the forward
        AND in the per-ray gradients; parameter gradients agree up to the fp32 summation order;
    (3) the backward is linear in the upstream gradient: doubling it doubles the per-ray gradients bit for bit (power-of-two
        scaling commutes with every bf16 / fp32 rounding in the chain) and the atomically accumulated ones (origin,
        parameters) up to the fp32 summation order;
    (4) the oracle agrees on a 96-ray slice of the same batch."""
    from model import _cabi, _ops
    R = 4096
    params, o, d, noise = _full_size_inputs(R)
    packed = _ops.pack_weights([p.detach() for p in params])
    rc = _ops.make_cfg(R, 128, 0.01, 10.0, 1, 0, 0, 1, _cabi.PREC_BF16)
    g = torch.Generator().manual_seed(1)
    p_rgb, p_dist = torch.randn(R, 3, generator=g).to(dev()) / R, torch.randn(R, generator=g).to(dev()) / R

    def run(perm=None, scale=1.0):
        for t in params + [o, d]:
            t.grad = None
        dd, nn, pr, pd = (d, noise, p_rgb, p_dist) if perm is None else (d[perm], noise[perm], p_rgb[perm], p_dist[perm])
        rgb, dist, w, a, z = _ops.RenderRays.apply(rc, packed, o, dd, nn, *params)
        (((rgb * pr).sum() + (dist * pd).sum()) * scale).backward()
        return rgb.detach(), dist.detach(), w.detach(), d.grad.clone(), o.grad.clone(), [p.grad.clone() for p in params]

    rgb, dist, w, dd, do, dp = run()
    assert float(rgb.min()) >= 0.0 and float(rgb.max()) <= 1.0 and float(w.min()) >= 0.0
    ws = w.sum(1)
    assert float(ws.max()) <= 1.0 + 1e-4
    ok = ws > 1e-3
    mean_z = dist[ok] / ws[ok]
    assert float(mean_z.min()) >= 0.01 - 1e-4 and float(mean_z.max()) <= 10.0 + 1e-3
    # (2) permutation equivariance
    perm = torch.randperm(R, generator=g).to(dev())
    rgb_p, dist_p, w_p, dd_p, do_p, dp_p = run(perm)
    assert torch.equal(rgb_p, rgb[perm]) and torch.equal(dist_p, dist[perm]) and torch.equal(w_p, w[perm])
    assert torch.equal(dd_p, dd)   # d.grad is scattered back through the index: same rows, same bits
    for a, b in zip(dp, dp_p):
        assert max_err(b, a) <= 5e-5 * float(a.abs().max()) + 1e-9   # fp32 sums over 524,288 samples in another order
    # (3) linearity in the upstream gradient
    _, _, _, dd2, do2, dp2 = run(scale=2.0)
    assert torch.equal(dd2, 2 * dd)              # per-ray gradients: no cross-CTA accumulation, bit for bit
    assert max_err(do2, 2 * do) <= 1e-5 * float(do.abs().max())   # atomically accumulated across CTAs: order varies
    for a, b in zip(dp, dp2):
        assert max_err(b, 2 * a) <= 2e-5 * float(a.abs().max()) + 1e-9
    # (4) oracle on a slice
    from oracle import nope_oracle as O
    sl = slice(1000, 1096)
    sd = syn.init_mlp_state(42)
    with torch.no_grad():
        zc = O.stratified_z(96, 128, 0.01, 10.0, noise[sl].cpu())
        oc, dc = o.detach().cpu(), d.detach().cpu()[sl]
        pts = (oc.reshape(1, 1, 3) + dc.unsqueeze(1) * zc.unsqueeze(-1)).reshape(-1, 3)
        view = (-dc).unsqueeze(1).expand(96, 128, 3).reshape(-1, 3)
        rgb_s, alpha = O.mlp_forward(sd, pts, view)
        rgb_o, dist_o, _, _ = O.composite(alpha.reshape(96, 128), rgb_s.reshape(96, 128, 3), zc)
    assert max_err(rgb[sl], rgb_o) <= 1e-2 and max_err(dist[sl], dist_o) <= 1e-1


def test_tc_render_chunking_invariance_cfg4_slice():
    """Full-frame rendering (cfg4) walks the frame in chunks: the result must not depend on where the chunk boundaries fall
    (odd chunk sizes exercise the half-filled last ray pair of a CTA)."""
    from model import _cabi, _ops
    R = 20001
    params, o, d, _ = _full_size_inputs(R, seed=9)
    params = [p.detach() for p in params]
    o, d = o.detach(), d.detach()
    packed = _ops.pack_weights(params)

    def render(chunk):
        outs = []
        for a in range(0, R, chunk):
            n = min(chunk, R - a)
            rc = _ops.make_cfg(n, 128, 0.01, 10.0, 1, 0, 0, 1, _cabi.PREC_BF16)
            rgb, dist, _, _, _ = _ops.render_nograd(rc, params, packed, o, d[a:a + n].contiguous(), None, want_aux=False)
            outs.append(torch.cat([rgb, dist.reshape(-1, 1)], 1))
        return torch.cat(outs)

    whole = render(R)
    assert torch.equal(render(4097), whole) and torch.equal(render(333), whole)


@pytest.mark.parametrize("precision", ["bf16", "fp32"])
@pytest.mark.parametrize("dist_alpha", [False, True])
def test_per_ray_origins_and_view_dirs_vs_oracle(dist_alpha, precision):
    """nnb_render_fwd_ex / nnb_render_bwd_ex (nnb_ray_extras: the NDC branch's per-ray origins and separate view directions),
    with and without dist_alpha (model/rendering.py:116-122), forward and every gradient, against the oracle
    (bf16 path: the bf16-emulating oracle, 5e-2 relative L2; fp32 path: the fp32 oracle, 2e-3)."""
    from model import _cabi, _ops
    from oracle import nope_oracle as O
    from oracle.bf16_emulation import mlp_forward_bf16emu
    R = 40
    g = torch.Generator().manual_seed(300 + int(dist_alpha))
    ro = torch.randn(R, 3, generator=g) * 0.3
    d = torch.nn.functional.normalize(torch.randn(R, 3, generator=g) + torch.tensor([0., 0., -2.]), dim=-1)
    vw = torch.nn.functional.normalize(torch.randn(R, 3, generator=g), dim=-1)
    noise = torch.rand(R, 128, generator=g)
    sd = syn.init_mlp_state(42)
    p_rgb, p_dist = torch.randn(R, 3, generator=g) / R, torch.randn(R, generator=g) / R
    mlp = mlp_forward_bf16emu if precision == "bf16" else O.mlp_forward

    ro_, d_, vw_ = ro.clone().requires_grad_(True), d.clone().requires_grad_(True), vw.clone().requires_grad_(True)
    P = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
    z = O.stratified_z(R, 128, 0.01, 10.0, noise)
    pts = (ro_.unsqueeze(1) + d_.unsqueeze(1) * z.unsqueeze(-1)).reshape(-1, 3)
    view = (-vw_).unsqueeze(1).expand(R, 128, 3).reshape(-1, 3)
    rgb_s, dens = mlp(P, pts, view, "softplus", dist_alpha)
    rgb, dist, w, alpha = O.composite(dens.reshape(R, 128), rgb_s.reshape(R, 128, 3), z, dist_alpha=dist_alpha)
    ((rgb * p_rgb).sum() + (dist * p_dist).sum()).backward()

    prec = _cabi.PREC_BF16 if precision == "bf16" else _cabi.PREC_FP32
    rc = _ops.make_cfg(R, 128, 0.01, 10.0, 1, int(dist_alpha), 0, 1, prec)
    params = [sd[k].to(dev()).requires_grad_(True) for k in _cabi.PARAM_ORDER]
    packed = _ops.pack_weights([p.detach() for p in params]) if precision == "bf16" else None
    rod, dd, vd = ro.to(dev()).requires_grad_(True), d.to(dev()).requires_grad_(True), vw.to(dev()).requires_grad_(True)
    rgb2, dist2, w2, a2, z2 = _ops.RenderRaysEx.apply(rc, packed, rod, dd, vd, noise.to(dev()), *params)
    ((rgb2 * p_rgb.to(dev())).sum() + (dist2 * p_dist.to(dev())).sum()).backward()
    torch.cuda.synchronize()
    assert torch.equal(z2.cpu(), z)
    ftol = 1e-2 if precision == "bf16" else 2e-4
    errs = dict(rgb=max_err(rgb2, rgb), alpha=max_err(a2, alpha), weights=max_err(w2, w), dist=max_err(dist2, dist) / 10.0)
    print("extras fwd dist_alpha=%s %s: %s" % (dist_alpha, precision, {k: "%.2e" % e for k, e in errs.items()}))
    assert max(errs.values()) <= ftol, errs
    gtol = 5e-2 if precision == "bf16" else 2e-3
    gerr = {k: _rel_l2(p.grad, P[k].grad) for k, p in zip(_cabi.PARAM_ORDER, params)}
    gerr["d_ray_o"], gerr["d_dirs"], gerr["d_view"] = _rel_l2(rod.grad, ro_.grad), _rel_l2(dd.grad, d_.grad), _rel_l2(vd.grad, vw_.grad)
    print("extras bwd dist_alpha=%s %s: worst rel-L2 %.3e" % (dist_alpha, precision, max(gerr.values())), {k: "%.1e" % e for k, e in gerr.items()})
    assert max(gerr.values()) <= gtol, gerr
    # forward-only entry point gives the same numbers
    rgb3, dist3, _, _, _ = _ops.render_nograd(rc, [p.detach() for p in params], packed, rod.detach(), dd.detach(), noise.to(dev()),
                                              ray_o=rod.detach(), view=vd.detach())
    assert max_err(rgb3, rgb2) <= (1e-6 if precision == "fp32" else 0.0) and max_err(dist3, dist2) <= (1e-5 if precision == "fp32" else 0.0)
