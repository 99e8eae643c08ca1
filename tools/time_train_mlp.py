"""Times the fused render forward(save) + backward (dgrad chain + wgrad) for R rays with CUDA events, per kernel."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "my-nope-nerf_b200")):
    sys.path.insert(0, p)
import torch
import synthetic_scene as syn
from model import _cabi, _ops

R = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
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


def step():
    for p in params:
        p.grad = None
    rgb, dist, w, a, z = _ops.RenderRays.apply(rc, packed, o, d, noise, *params)
    ((rgb * p_rgb).sum() + dist.sum() / R).backward()


for _ in range(3):
    step()
torch.cuda.synchronize()
_cabi.profile = {}
for _ in range(iters):
    step()
_cabi.profile_flush()
prof = {k: v[0] / iters for k, v in _cabi.profile.items()}
_cabi.profile = None
tot = prof.get("nnb_render_fwd", 0) + prof.get("nnb_render_bwd", 0) + prof.get("nnb_render_bwd_ex", 0)
print(json.dumps({"rays": R, "ms": prof, "mlp_ms": tot, "tflops": 455_737_344 * R / tot / 1e9}))
