"""Times the geometry render (sphere tracing + phong shading, model/rendering.py:203-385) of R pixels with CUDA events:
fp32 point kernels for all 512 proposals per ray vs the fused tensor-core kernel for the proposals (bf16 renderer)."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "my-nope-nerf_b200")):
    sys.path.insert(0, p)
import torch
import synthetic_scene as syn
import model as mdl
from model import geometry

R = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
dev = torch.device("cuda:0")
cfg = syn.default_cfg()
sd = {k: v.clone() for k, v in syn.init_mlp_state(42).items()}
sd["fc_density.weight"] = sd["fc_density.weight"] * 200.0
sd["fc_density.bias"] = torch.tensor([0.86])
nerf = mdl.OfficialStaticNerf(cfg)
nerf.load_state_dict(sd)
nerf = nerf.to(dev)
g = torch.Generator().manual_seed(5)
cam = torch.tensor([0.1, -0.2, -0.9], device=dev)
dirs = torch.nn.functional.normalize(torch.randn(R, 3, generator=g) * torch.tensor([0.4, 0.4, 0.1]) + torch.tensor([0., 0., 1.]), dim=-1).to(dev)
out = {"pixels": R}
for prec in ("fp32", "bf16"):
    r = mdl.Renderer(nerf, dict(cfg["rendering"], precision=prec), device=dev)
    ts = []
    for i in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        o = geometry.phong_render(r, cam, dirs, 0)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    out[prec + "_ms"] = min(ts)
    out[prec + "_hit_frac"] = float((o["rgb_surf"].abs().sum(-1) > 0).float().mean())
print(json.dumps(out))
