"""Experiment (round 2): can the Chamfer search hide beside the HBM-bound weight-gradient kernel?
Times, at cfg2 sizes (4,096 rays; 28,830-point clouds, both directions):
  A. render backward alone (tc_bwd then tc_wgrad on its own stream)
  B. Chamfer alone, tcgen05 sweeps (production) and CUDA-core sweeps (no TMEM, 16 KB smem per 128-thread block)
  C. render backward with the CUDA-core Chamfer launched on a third stream once tc_bwd has finished (i.e. beside tc_wgrad)
  D. the same with the tcgen05 Chamfer (cannot co-reside: wgrad owns all 512 TMEM columns) for comparison"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "my-nope-nerf_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402
import synthetic_scene as syn  # noqa: E402
from model import _cabi, _ops  # noqa: E402

R = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
dev = torch.device("cuda:0")
sd = syn.init_mlp_state(42)
params = [sd[k].to(dev).requires_grad_(True) for k in _cabi.PARAM_ORDER]
g = torch.Generator().manual_seed(1)
o = torch.zeros(3, device=dev, requires_grad=True)
d = torch.nn.functional.normalize(torch.randn(R, 3, generator=g) + torch.tensor([0., 0., -2.]), dim=-1).to(dev).requires_grad_(True)
noise = torch.rand(R, 128, generator=g).to(dev)
packed = _ops.pack_weights([p.detach() for p in params])
rc = _ops.make_cfg(R, 128, 0.01, 10.0, 1, 0, 0, 1, _cabi.PREC_BF16)
p_rgb = torch.randn(R, 3, generator=g).to(dev) / R
P = 28830
X = (torch.rand(P, 3, generator=g) * 2 - 1).to(dev)
Y = (X + 0.07 * torch.randn(P, 3, generator=g).to(dev)).contiguous()    # ~9 point spacings apart, like the bench's random poses
wg, cs = torch.cuda.Stream(), torch.cuda.Stream()
L = _cabi.lib()


def chamfer(mode):
    old = L.nnb_debug_chamfer_pass1(mode)
    _ops.chamfer_nn(X, Y)
    _ops.chamfer_nn(Y, X)
    L.nnb_debug_chamfer_pass1(old)


def run(with_chamfer, mode, beside):
    """-> ms from the start of the backward to everything done"""
    for p in params:
        p.grad = None
    rgb, dist, w, a, z = _ops.RenderRays.apply(rc, packed, o, d, noise, *params)
    loss = (rgb * p_rgb).sum() + dist.sum() / R
    torch.cuda.synchronize()
    main = torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    _ops.wgrad_stream, _ops.wgrad_forked = wg, False
    e0.record()
    loss.backward()                      # tc_bwd on main, tc_wgrad forked onto wg
    if with_chamfer:
        if beside:
            cs.wait_stream(main)         # starts when tc_bwd (and the small tail behind it) is done: beside tc_wgrad
            with torch.cuda.stream(cs):
                chamfer(mode)
            main.wait_stream(cs)
        else:
            main.wait_stream(wg)
            chamfer(mode)
    main.wait_stream(wg)
    _ops.wgrad_stream = None
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


def med(f, n=7):
    f()
    f()
    v = sorted(f() for _ in range(n))
    return v[len(v) // 2]


def alone(mode):
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    chamfer(mode)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1)


out = {"rays": R, "points": P,
       "A_backward_alone_ms": med(lambda: run(False, 0, False)),
       "B_chamfer_alone_tcgen05_ms": med(lambda: alone(0)), "B_chamfer_alone_cuda_cores_ms": med(lambda: alone(1)),
       "serial_backward_then_tcgen05_ms": med(lambda: run(True, 0, False)),
       "C_backward_with_cuda_core_chamfer_beside_wgrad_ms": med(lambda: run(True, 1, True)),
       "D_backward_with_tcgen05_chamfer_beside_wgrad_ms": med(lambda: run(True, 0, True))}
print(json.dumps(out, indent=1))
