"""Shared helpers for the parity tests."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "my-nope-nerf_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

import synthetic_scene as syn  # noqa: E402


def dev():
    return torch.device("cuda:0")


def to_dev(x):
    if torch.is_tensor(x):
        return x.to(dev())
    if isinstance(x, dict):
        return {k: to_dev(v) for k, v in x.items()}
    return x


def max_err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max())


def rel_err(a, b):
    """max |a-b| / max|b| : scale-relative error of a whole tensor."""
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


def assert_close(a, b, rtol, atol, name=""):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    assert a.shape == b.shape, "%s shape %s vs %s" % (name, tuple(a.shape), tuple(b.shape))
    diff = (a - b).abs()
    tol = atol + rtol * b.abs()
    bad = diff > tol
    if bad.any():
        i = torch.nonzero(bad.reshape(-1))[0].item()
        raise AssertionError("%s: %d/%d elements out of tolerance (rtol %g atol %g); max abs err %.3e, first bad: got %r want %r"
                             % (name, int(bad.sum()), bad.numel(), rtol, atol, float(diff.max()),
                                float(a.reshape(-1)[i]), float(b.reshape(-1)[i])))


def build_models(cfg, n_cams, scene=None, precision="fp32", state=None, r0=None, t0=None, scales=None, shifts=None):
    """The drop-in module stack on cuda:0 with the given parameter values."""
    import model as mdl
    d = dev()
    nerf = mdl.OfficialStaticNerf(cfg)
    nerf.load_state_dict(state if state is not None else syn.init_mlp_state(42))
    rcfg = dict(cfg["rendering"])
    rcfg["precision"] = precision
    renderer = mdl.Renderer(nerf, rcfg, device=d)
    net = mdl.get_model(renderer, cfg, device=d).to(d)
    pose = mdl.LearnPose(n_cams, True, True, cfg, init_c2w=None if scene is None else scene["poses_gt"].clone()).to(d)
    dist = mdl.Learn_Distortion(n_cams, True, True, cfg).to(d)
    with torch.no_grad():
        if r0 is not None:
            pose.r.copy_(r0)
            pose.t.copy_(t0)
        if scales is not None:
            dist.global_scales.copy_(scales)
            dist.global_shifts.copy_(shifts)
    return nerf, renderer, net, pose, dist
