"""Memory-safety checks that need no external tool (compute-sanitizer is closed on this pool: it exits 86 with "closed on this pool
... find a bad access with bounds checks and asserts of your own"; profiles/r2_sanitizer.txt).

Every tensor the autograd shims of model/_ops.py allocate -- kernel outputs, gradients, saved-activation and Chamfer workspaces --
is placed between two 4 KB canary zones and pre-filled with 0xFF bytes (a NaN pattern for fp32, -1 for integers).  After a call:
  * a canary that changed          = an out-of-bounds WRITE (what memcheck reports),
  * a NaN / changed result         = a READ of memory the kernels never wrote, e.g. uninitialised workspace (what initcheck reports),
and the results must equal those of the ordinary run.  Races are approached by repeatability: kernels without floating-point
atomics must return bit-identical results when repeated."""
import pytest
import torch

from util import build_models, dev, syn

pytestmark = pytest.mark.gpu

GUARD = 4096


class _GuardedTorch(object):
    """Stands in for the `torch` module inside model._ops: the allocating factories hand out guarded, poisoned buffers."""

    def __init__(self):
        self.__dict__["_allocs"] = []

    def __getattr__(self, name):
        return getattr(torch, name)

    def _alloc(self, shape, dtype, device, fill=None):
        if len(shape) == 1 and isinstance(shape[0], (tuple, list, torch.Size)):
            shape = tuple(shape[0])
        n = 1
        for d in shape:
            n *= int(d)
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        pad = (-nbytes) % 256
        raw = torch.empty(GUARD + nbytes + pad + GUARD, dtype=torch.uint8, device=device)
        raw[:GUARD] = 0xA5
        raw[GUARD + nbytes:] = 0xA5
        payload = raw[GUARD:GUARD + nbytes]
        payload.fill_(0xFF if fill is None else 0)
        self._allocs.append((raw, nbytes))
        t = payload.view(dtype).view(*shape) if n else torch.empty(shape, dtype=dtype, device=device)
        if fill not in (None, 0):
            t.fill_(fill)
        return t

    def empty(self, *shape, dtype=torch.float32, device=None, **kw):
        return self._alloc(shape, dtype, device)

    def zeros(self, *shape, dtype=torch.float32, device=None, **kw):
        return self._alloc(shape, dtype, device, fill=0)

    def empty_like(self, t, **kw):
        return self._alloc(tuple(t.shape), kw.get("dtype", t.dtype), t.device)

    def zeros_like(self, t, **kw):
        return self._alloc(tuple(t.shape), kw.get("dtype", t.dtype), t.device, fill=0)

    def check(self):
        torch.cuda.synchronize()
        bad = 0
        for raw, nbytes in self._allocs:
            if not bool((raw[:GUARD] == 0xA5).all()) or not bool((raw[GUARD + nbytes:] == 0xA5).all()):
                bad += 1
        return len(self._allocs), bad


@pytest.fixture
def guarded(monkeypatch):
    from model import _ops
    g = _GuardedTorch()
    monkeypatch.setattr(_ops, "torch", g)
    monkeypatch.setattr(_ops, "_ws_cache", {})        # cached workspaces are re-allocated under guard
    yield g


def _step(precision, depth_loss_type="l1"):
    import model as mdl
    cfg = syn.default_cfg()
    scene = syn.make_scene("tiny")
    n, R = scene["n"], 61           # odd ray count: the last CTA pair of the tensor-core kernels runs a dummy tile
    r0, t0 = syn.init_pose_params(n)
    nerf, renderer, net, pose, dist = build_models(cfg, n, scene, precision=precision, r0=r0, t0=t0)
    tcfg = dict(cfg["training"], n_training_points=R, depth_loss_type=depth_loss_type)
    tr = mdl.Trainer(net, None, tcfg, device=dev(), pose_param_net=pose, distortion_net=dist)
    tr.fixed_ray_idx, tr.fixed_noise = syn.ray_choice(scene["h"], scene["w"], R), syn.jitter(R)
    losses, grads = [], []
    for idx, ref in ((1, 2), (n - 1, n - 2)):
        for p in list(nerf.parameters()) + list(pose.parameters()) + list(dist.parameters()):
            p.grad = None
        ld = tr.compute_loss(syn.make_batch(scene, idx, ref), it=1, epoch=0, scheduling_start=cfg["training"]["scheduling_start"])
        ld["loss"].backward()
        torch.cuda.synchronize()
        losses.append({k: float(ld[k].detach()) for k in ("loss", "loss_rgb", "loss_depth", "loss_pc", "loss_rgb_s")})
        grads.append([p.grad.detach().clone() for p in list(nerf.parameters()) + [pose.r, pose.t]])
    return losses, grads


@pytest.mark.parametrize("precision,dlt", [("fp32", "l1"), ("bf16", "l1"), ("fp32", "invariant")])
def test_training_step_writes_inside_its_buffers_and_reads_nothing_unwritten(guarded, monkeypatch, precision, dlt):
    from model import _ops
    got_l, got_g = _step(precision, dlt)
    n_alloc, bad = guarded.check()
    print("%s/%s: %d guarded allocations, %d damaged canaries" % (precision, dlt, n_alloc, bad))
    assert n_alloc > 40 and bad == 0
    monkeypatch.setattr(_ops, "torch", torch)           # the ordinary run, for comparison
    monkeypatch.setattr(_ops, "_ws_cache", {})
    want_l, want_g = _step(precision, dlt)
    for a, b in zip(got_l, want_l):
        for k in a:
            assert a[k] == a[k], "NaN loss %s: something read poisoned memory" % k
            assert abs(a[k] - b[k]) <= 1e-6 * max(1.0, abs(b[k])), (k, a[k], b[k])
    for ga, gb in zip(got_g, want_g):
        for x, y in zip(ga, gb):
            assert bool(torch.isfinite(x).all())
            # weight-gradient atomics add in another order: 1e-4 of the tensor's scale, with a floor for near-cancelling scalars
            assert float((x - y).abs().max()) <= 1e-4 * float(y.abs().max()) + 1e-7


@pytest.mark.parametrize("n,m", [(1, 1), (129, 257), (4099, 1023), (28830, 28829)])
@pytest.mark.parametrize("mode", [0, 1, 2])
def test_chamfer_writes_inside_its_buffers(guarded, monkeypatch, n, m, mode):
    from model import _ops, _cabi
    g = torch.Generator().manual_seed(n + m)
    src, dst = (torch.rand(n, 3, generator=g) * 2 - 1).to(dev()), (torch.rand(m, 3, generator=g) * 2 - 1).to(dev())
    old = _cabi.lib().nnb_debug_chamfer_pass1(mode)
    try:
        idx, dist = _ops.chamfer_nn(src, dst, want_dist=True)
        n_alloc, bad = guarded.check()
        assert n_alloc == 3 and bad == 0               # idx, dist, workspace
        monkeypatch.setattr(_ops, "torch", torch)
        monkeypatch.setattr(_ops, "_ws_cache", {})
        idx2, dist2 = _ops.chamfer_nn(src, dst, want_dist=True)
    finally:
        _cabi.lib().nnb_debug_chamfer_pass1(old)
    assert torch.equal(idx, idx2) and torch.equal(dist, dist2)
    assert int(idx.min()) >= 0 and int(idx.max()) < m


def test_full_frame_render_and_point_queries_write_inside_their_buffers(guarded):
    import model as mdl
    cfg = syn.default_cfg()
    scene = syn.make_scene("tiny")
    for precision in ("fp32", "bf16"):
        nerf, renderer, net, pose, dist = build_models(cfg, scene["n"], scene, precision=precision)
        ex = mdl.Extract_Images(renderer, {"extract_images": {"resolution": [scene["h"] - 1, scene["w"] - 3]}}, device=dev(), render_type="nope_nerf")
        with torch.no_grad():
            wm = torch.inverse(scene["poses_gt"][1]).unsqueeze(0).to(dev())
            rgb, depth = ex.render_frame(scene["camera_mat"].to(dev()), wm, scene["scale_mat"].to(dev()), scene["h"] - 1, scene["w"] - 3)
            pts = (torch.rand(1001, 3, generator=torch.Generator().manual_seed(2)) - 0.5).to(dev()) * 4
            grad = nerf.gradient(pts, 0)
        n_alloc, bad = guarded.check()
        assert bad == 0 and n_alloc >= 5
        assert bool(torch.isfinite(rgb).all()) and bool(torch.isfinite(depth).all()) and bool(torch.isfinite(grad).all())


def test_kernels_without_float_atomics_repeat_bit_identically():
    """A data race shows as run-to-run differences.  Forward render (both precisions), per-ray backward outputs, ray generation and
    the Chamfer search contain no floating-point atomics: ten repeats must agree bit for bit."""
    from model import _ops, _cabi
    cfg = syn.default_cfg()
    scene = syn.make_scene("tiny")
    g = torch.Generator().manual_seed(3)
    R = 333
    dirs = torch.nn.functional.normalize(torch.randn(R, 3, generator=g) + torch.tensor([0., 0., -2.]), dim=-1).to(dev())
    origin = torch.zeros(3, device=dev())
    noise = torch.rand(R, 128, generator=g).to(dev())
    sd = syn.init_mlp_state(42)
    params = [sd[k].to(dev()) for k in _cabi.PARAM_ORDER]
    src, dst = (torch.rand(5000, 3, generator=g)).to(dev()), (torch.rand(7001, 3, generator=g)).to(dev())
    for prec in (_cabi.PREC_FP32, _cabi.PREC_BF16):
        rc = _ops.make_cfg(R, 128, 0.01, 10.0, 1, 0, 0, 1, prec)
        packed = _ops.pack_weights(params) if prec == _cabi.PREC_BF16 else None
        first = None
        for rep in range(10):
            out = _ops.render_nograd(rc, params, packed, origin, dirs, noise)
            torch.cuda.synchronize()
            if first is None:
                first = [t.clone() for t in out]
            else:
                for a, b in zip(first, out):
                    assert torch.equal(a, b), "render forward differs between repeats (precision %d, repeat %d)" % (prec, rep)
    first = None
    for rep in range(10):
        idx, dist = _ops.chamfer_nn(src, dst, want_dist=True)
        if first is None:
            first = (idx.clone(), dist.clone())
        else:
            assert torch.equal(first[0], idx) and torch.equal(first[1], dist)
