"""CPU-side checks: the C-ABI library builds, loads and exports every symbol the header declares; host logic
(annealing, sharding, flat gradient buffers); drop-in surface parity with the reference (when it is present);
multi-process data-parallel plumbing over gloo (world size 2)."""
import ctypes
import inspect
import os
import re
import subprocess
import sys

import pytest
import torch

from util import PKG, ROOT, syn


def _header_functions():
    txt = open(os.path.join(ROOT, "include", "nope_nerf_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(nnb_[a-z0-9_]+)\s*\(", txt)))


def test_library_builds_and_exports_every_declared_symbol():
    sys.path.insert(0, PKG)
    import build as nnb_build
    so = nnb_build.build()
    assert os.path.exists(so)
    lib = ctypes.CDLL(so)
    names = _header_functions()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), "missing export %s" % n
    from model import _cabi
    assert sorted(_cabi.SIGNATURES.keys()) == names, set(names) ^ set(_cabi.SIGNATURES.keys())
    lib.nnb_version.restype = ctypes.c_int
    assert lib.nnb_version() >= 100
    lib.nnb_error_string.restype = ctypes.c_char_p
    assert b"workspace" in lib.nnb_error_string(-2)


def test_sass_is_sm100a():
    so = os.path.join(PKG, "libnope_nerf_b200.so")
    if not os.path.exists(so):
        pytest.skip("library not built")
    out = subprocess.run(["cuobjdump", "-lelf", so], capture_output=True, text=True).stdout
    assert "sm_100a" in out, out[:500]


def test_product_refuses_cpu_tensors():
    from model import _ops
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        _ops.RgbLoss.apply(torch.zeros(1, 4, 3), torch.zeros(1, 4, 3))


def test_product_never_imports_oracle():
    for root, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(root, f)).read()
                assert "nope_oracle" not in txt and "from oracle" not in txt and "import oracle" not in txt, f


def test_state_dict_layout():
    import model as mdl
    cfg = syn.default_cfg()
    nerf = mdl.OfficialStaticNerf(cfg)
    net = mdl.get_model(mdl.Renderer(nerf, cfg["rendering"], device="cpu"), cfg, device="cpu")
    sd = net.state_dict()
    want = {"renderer.model.layers0.0.weight": (256, 63), "renderer.model.layers1.0.weight": (256, 319),
            "renderer.model.fc_density.weight": (1, 256), "renderer.model.fc_feature.weight": (256, 256),
            "renderer.model.rgb_layers.0.weight": (128, 283), "renderer.model.fc_rgb.weight": (3, 128)}
    for k, shp in want.items():
        assert tuple(sd[k].shape) == shp
    assert len(sd) == 24 and sum(v.numel() for v in sd.values()) == 595844
    assert float(sd["renderer.model.fc_density.bias"]) == pytest.approx(0.1)
    assert len([m for m in net.modules() if isinstance(m, torch.nn.Linear)]) == 12  # train.py:347-350 'reset' mode
    pose = mdl.LearnPose(5, True, False, cfg, init_c2w=torch.eye(4).repeat(5, 1, 1))
    assert list(pose.state_dict().keys()) == ["init_c2w", "r", "t"] and not pose.t.requires_grad
    assert list(mdl.LearnPose(5, True, True, cfg).state_dict().keys()) == ["r", "t"]
    assert list(mdl.Learn_Distortion(5, True, True, cfg).state_dict().keys()) == ["global_scales", "global_shifts"]


def test_surface_matches_reference():
    """Constructor / forward signatures and exported names equal the reference's (only where it is mounted)."""
    from oracle import ref_harness
    if not ref_harness.reference_available():
        pytest.skip("reference tree not mounted")
    code = r'''
import sys, inspect, json
sys.path.insert(0, %r)
from oracle import ref_harness
ref = ref_harness.load_reference().model
names = ["OfficialStaticNerf", "Renderer", "nope_nerf", "LearnPose", "Learn_Distortion", "LearnFocal", "Trainer", "Trainer_pose", "CheckpointIO"]
out = {}
for n in names:
    c = getattr(ref, n)
    out[n] = {m: list(inspect.signature(getattr(c, m)).parameters) for m in ("__init__", "forward", "train_step", "compute_loss", "save", "load", "get_t")
              if hasattr(c, m) and callable(getattr(c, m))}
out["get_model"] = list(inspect.signature(ref.get_model).parameters)
import yaml
cfg = yaml.safe_load(open(ref_harness.REFERENCE_ROOT + "/configs/default.yaml"))
import torch
torch.manual_seed(42)          # train.py:24, then the construction order of train.py:47-56,97,110
net = ref.get_model(ref.Renderer(ref.OfficialStaticNerf(cfg), cfg["rendering"], device="cpu"), cfg, device="cpu")
pose = ref.LearnPose(5, True, True, cfg, init_c2w=None)
dist = ref.Learn_Distortion(5, True, True, cfg)
after = float(torch.rand(1))
out["state"] = {k: list(v.shape) for k, v in net.state_dict().items()}
out["seeded"] = {"net": {k: [float(v.double().sum()), float(v.double().abs().sum())] for k, v in net.state_dict().items()},
                 "pose": {k: float(v.double().abs().sum()) for k, v in pose.state_dict().items()},
                 "dist": {k: float(v.double().sum()) for k, v in dist.state_dict().items()}, "next_draw": after}
print(json.dumps(out))
''' % ROOT
    import json
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    ref = json.loads(res.stdout.strip().splitlines()[-1])
    import model as mdl
    for n, methods in ref.items():
        if n in ("get_model", "state", "seeded"):
            continue
        c = getattr(mdl, n)
        for m, params in methods.items():
            mine = list(inspect.signature(getattr(c, m)).parameters)
            assert mine[:len(params)] == params or set(params) <= set(mine), (n, m, params, mine)
    assert list(inspect.signature(mdl.get_model).parameters) == ref["get_model"]
    cfg = syn.default_cfg()
    net = mdl.get_model(mdl.Renderer(mdl.OfficialStaticNerf(cfg), cfg["rendering"], device="cpu"), cfg, device="cpu")
    assert {k: list(v.shape) for k, v in net.state_dict().items()} == ref["state"]
    # the same seed gives the same initial weights AND leaves the generator where the reference leaves it: a run of the unchanged
    # train.py (torch.manual_seed(42), train.py:24) starts from the reference's own initialisation
    torch.manual_seed(42)
    net = mdl.get_model(mdl.Renderer(mdl.OfficialStaticNerf(cfg), cfg["rendering"], device="cpu"), cfg, device="cpu")
    pose = mdl.LearnPose(5, True, True, cfg, init_c2w=None)
    dist = mdl.Learn_Distortion(5, True, True, cfg)
    after = float(torch.rand(1))
    sd = ref["seeded"]
    assert {k: [float(v.double().sum()), float(v.double().abs().sum())] for k, v in net.state_dict().items()} == sd["net"]
    assert {k: float(v.double().abs().sum()) for k, v in pose.state_dict().items()} == sd["pose"]
    assert {k: float(v.double().sum()) for k, v in dist.state_dict().items()} == sd["dist"] and after == sd["next_draw"]


def test_anneal_and_weights_match_oracle():
    import model as mdl
    from oracle import nope_oracle as O
    cfg = syn.default_cfg()["training"]
    tr = mdl.Trainer(None, None, cfg, device="cpu")
    for epoch in (0, 9999, 10000, 10500, 11999, 12000, 20000):
        w, kind = O.step_weights(cfg, epoch, cfg["scheduling_start"])
        for k in w:
            assert tr.anneal(cfg[k][0], cfg[k][1], cfg["scheduling_start"], cfg["annealing_epochs"], epoch) == w[k]


def test_shard_ranges_partition():
    from model.parallel import shard_range
    for n in (0, 1, 7, 4096, 65536, 465750):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_range(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1


def _dp_loss_forward_check(dp, world, rank, _ops, losses):
    """Loss.forward under data parallelism on gloo, the loss kernels replaced by torch stand-ins of the same
    signature: per-rank terms are G x (share of the GLOBAL normaliser), the invariant depth loss all-gathers the rendered depths
    (parallel.AllGatherRows), and the loss scalars ride in the gradient allreduce.  Unequal shards (61 rays) and a ragged mask: the
    rank mean of values and gradients must be the one-rank result."""
    import torch.distributed as dist
    from model import parallel

    class _Rgb(torch.autograd.Function):      # _ops.RgbLoss: [l1_sum / R, l2_sum / R, mse]
        @staticmethod
        def forward(ctx, pred, gt):
            ctx.save_for_backward(pred, gt)
            d = (pred - gt).reshape(-1, 3)
            return torch.stack([d.abs().sum() / d.shape[0], (d * d).sum() / d.shape[0], (d * d).mean()])

        @staticmethod
        def backward(ctx, g):
            pred, gt = ctx.saved_tensors
            d = pred - gt
            R = d.reshape(-1, 3).shape[0]
            return g[0] * torch.sign(d) / R + (g[1] / R + g[2] / (3 * R)) * 2 * d, None

    def depth_l1(pred, gt, mask):
        m = mask if mask is not None else torch.ones_like(pred, dtype=torch.bool)
        return ((pred - gt).abs() * m).sum() / m.sum().clamp_min(1)

    def depth_inv(pred, gt, mask):
        p, g_ = pred[mask], gt[mask]
        n = lambda x: (x - x.median()) / (x - x.median()).abs().mean()
        return ((n(p) - n(g_)) ** 2).mean()

    _ops.RgbLoss = _Rgb
    _ops.DepthL1 = type("L1", (), {"apply": staticmethod(depth_l1)})
    _ops.DepthInvariant = type("Inv", (), {"apply": staticmethod(depth_inv)})
    g = torch.Generator().manual_seed(11)
    R = 61
    pred0, gt0 = torch.rand(1, R, 3, generator=g), torch.rand(1, R, 3, generator=g)
    dpred0, dgt0 = torch.rand(R, generator=g) + 0.5, torch.rand(R, generator=g) + 0.5
    mask = torch.rand(R, generator=g) > 0.3
    w = {k: 0.0 for k in syn.WEIGHT_NAMES}
    w.update(rgb_weight=1.0, depth_weight=0.04)
    for dlt in ("l1", "invariant"):
        cfg = dict(syn.default_cfg()["training"], depth_loss_type=dlt)
        one = losses.Loss(cfg)
        a, b, c = pred0.clone().requires_grad_(True), dpred0.clone().requires_grad_(True), dgt0.clone().requires_grad_(True)
        ld1 = one(a, gt0, b, c, weights=w, rgb_loss_type="l1", depth_mask=mask)
        ld1["loss"].backward()
        L = losses.Loss(cfg)
        L.dp = dp
        lo, hi = dp.shard(R)
        pa, pb = pred0[:, lo:hi].clone().requires_grad_(True), dpred0[lo:hi].clone().requires_grad_(True)
        gall = dgt0.clone().requires_grad_(True)
        info = {"lo": lo, "hi": hi, "n_global": R, "mask_global": mask, "depth_gt_global": gall}
        ld = L(pa, gt0[:, lo:hi], pb, gall[lo:hi], weights=w, rgb_loss_type="l1", depth_mask=mask[lo:hi], dp_info=info)
        ld["loss"].backward()
        # what Trainer does after the backward: ONE allreduce over [gradients | loss scalars]
        gp = [torch.nn.Parameter(torch.zeros(1, R, 3)), torch.nn.Parameter(torch.zeros(R)), torch.nn.Parameter(torch.zeros(R))]
        gp[0].grad = torch.zeros(1, R, 3)
        gp[0].grad[:, lo:hi] = pa.grad
        gp[1].grad = torch.zeros(R)
        gp[1].grad[lo:hi] = pb.grad
        gp[2].grad = gall.grad if gall.grad is not None else torch.zeros(R)
        scal = [ld[k] for k in ("loss", "loss_rgb", "loss_depth", "l2_mean")]
        dp.allreduce_grads(gp, extra=scal)
        for k in ("loss", "loss_rgb", "loss_depth", "l2_mean"):
            assert abs(float(ld[k]) - float(ld1[k])) <= 2e-6 * max(1.0, abs(float(ld1[k]))), (dlt, k, float(ld[k]), float(ld1[k]))
        for got, want in ((gp[0].grad, a.grad), (gp[1].grad, b.grad), (gp[2].grad, c.grad)):
            assert torch.allclose(got, want, rtol=1e-4, atol=1e-7), (dlt, float((got - want).abs().max()))


def _gloo_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    sys.path.insert(0, PKG)
    import torch.distributed as dist
    from model import parallel
    dp = parallel.init_from_env(backend="gloo")
    assert dp.world_size == world and dp.rank == rank
    # every rank draws the same global batch and takes its shard: the union is the 1-rank batch
    idx = syn.ray_choice(24, 32, 100, seed=42)
    lo, hi = dp.shard(idx.numel())
    got = dp.gather_rows(idx[lo:hi].reshape(-1, 1).float(), idx.numel()).reshape(-1).long()
    assert torch.equal(got, idx)
    # gradient averaging over one flat buffer; a parameter the step did not reach (no gradient on any rank) keeps none, so
    # that the optimiser skips it as torch.optim.Adam does under the reference (tests/test_gpu_trajectory.py)
    torch.manual_seed(0)
    ps = [torch.nn.Parameter(torch.randn(3, 5)), torch.nn.Parameter(torch.randn(7)), torch.nn.Parameter(torch.randn(2))]
    ps[0].grad = torch.full((3, 5), float(rank + 1))
    ps[1].grad = torch.arange(7.0) * (rank + 1)
    dp.check_symmetry = True
    dp.allreduce_grads(ps)
    mean = sum(r + 1 for r in range(world)) / world
    assert torch.allclose(ps[0].grad, torch.full((3, 5), mean))
    assert torch.allclose(ps[1].grad, torch.arange(7.0) * mean)
    assert ps[2].grad is None
    # ranks that disagree on which parameters have a gradient are a bug upstream: the check names the parameter
    if rank == 0:
        ps[2].grad = torch.ones(2)
    try:
        dp.allreduce_grads(ps)
        raise AssertionError("asymmetric gradients not reported")
    except RuntimeError as e:
        assert "[2]" in str(e)
    ps[2].grad = None
    dp.check_symmetry = False
    # Chamfer with the SOURCE points sharded by rank: averaging the ranks' terms and gradients must give the
    # one-rank loss and gradient.  The two kernels are replaced by torch stand-ins: this checks the host-side sharding.
    from model import _ops, losses

    class _Err(torch.autograd.Function):  # signature of _ops.ChamferErr
        @staticmethod
        def forward(ctx, X, Y, idx):
            ctx.save_for_backward(X, Y, idx)
            return (X - Y[idx]).norm(dim=1).mean()

        @staticmethod
        def backward(ctx, g):
            X, Y, idx = ctx.saved_tensors
            d = X - Y[idx]
            u = d / d.norm(dim=1, keepdim=True).clamp_min(1e-30) * (g / X.shape[0])
            return u, torch.zeros_like(Y).index_add_(0, idx, -u), None

    _ops.chamfer_nn = lambda src, dst, want_dist=False: torch.cdist(src.detach(), dst.detach()).argmin(dim=1)
    _ops.ChamferErr = _Err
    g = torch.Generator().manual_seed(5)
    X0, Y0 = torch.rand(1, 301, 3, generator=g), torch.rand(1, 257, 3, generator=g)
    L = losses.Loss(syn.default_cfg()["training"])
    Xf, Yf = X0.clone().requires_grad_(True), Y0.clone().requires_grad_(True)
    full = L.get_pc_loss(Xf, Yf)
    full.backward()
    L.dp = dp
    Xs, Ys = X0.clone().requires_grad_(True), Y0.clone().requires_grad_(True)
    part = L.get_pc_loss(Xs, Ys)
    part.backward()
    red = torch.cat([part.detach().reshape(1), Xs.grad.reshape(-1), Ys.grad.reshape(-1)])
    dist.all_reduce(red)
    red /= world
    want = torch.cat([full.detach().reshape(1), Xf.grad.reshape(-1), Yf.grad.reshape(-1)])
    assert torch.allclose(red, want, rtol=1e-5, atol=1e-7), float((red - want).abs().max())
    _dp_loss_forward_check(dp, world, rank, _ops, losses)
    open(os.path.join(out_dir, "ok%d" % rank), "w").write("ok")
    dist.destroy_process_group()


def test_data_parallel_plumbing_gloo_world2(tmp_path):
    import torch.multiprocessing as mp
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_gloo_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(tmp_path / "ok0") and os.path.exists(tmp_path / "ok1")


def test_resident_scene_schema_on_cpu():
    """ResidentScene is plain tensor bookkeeping: on the CPU device it must reproduce the batch schema and the
    reference-frame rule of dataloading/dataset.py:295-301."""
    import random
    sys.path.insert(0, PKG)
    from model.scene import ResidentScene
    scene = syn.make_scene("tiny")
    n = scene["n"]
    rs = ResidentScene([syn.make_batch(scene, i) for i in range(n)], torch.device("cpu"), random_ref=2)
    rng = random.Random(0)
    for f in range(n):
        for _ in range(4):
            r = rs.ref_index(f, rng)
            assert (r == f - 1) if f == n - 1 else (f + 1 <= r <= min(f + 2, n - 1))
    b, hb = rs.batch(2, 3), syn.make_batch(scene, 2, 3)
    assert set(b) == set(hb)
    for k, v in hb.items():
        assert torch.equal(b[k].reshape(v.shape), v), k
    with pytest.raises(ValueError):
        ResidentScene([syn.make_batch(scene, 0), syn.make_batch(scene, 2)], torch.device("cpu"))
