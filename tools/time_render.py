"""Times the fused render forward (one full 1242x375 frame = 465,750 rays x 128 samples) with CUDA events."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "my-nope-nerf_b200")):
    sys.path.insert(0, p)
import torch
import synthetic_scene as syn
from model import _cabi, _ops

R = int(sys.argv[1]) if len(sys.argv) > 1 else 465750
prec = sys.argv[2] if len(sys.argv) > 2 else "bf16"
dev = torch.device("cuda:0")
sd = syn.init_mlp_state(42)
params = [sd[k].to(dev) for k in _cabi.PARAM_ORDER]
g = torch.Generator().manual_seed(1)
o = torch.zeros(3, device=dev)
d = torch.nn.functional.normalize(torch.randn(R, 3, generator=g) + torch.tensor([0., 0., -2.]), dim=-1).to(dev)
P = _cabi.PREC_BF16 if prec == "bf16" else _cabi.PREC_FP32
packed = _ops.pack_weights(params) if prec == "bf16" else None
rc = _ops.make_cfg(R, 128, 0.01, 10.0, 1, 0, 0, 1, P)
for _ in range(3):
    _ops.render_nograd(rc, params, packed, o, d, None, want_aux=False)
torch.cuda.synchronize()
ts = []
for _ in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    _ops.render_nograd(rc, params, packed, o, d, None, want_aux=False)
    e1.record()
    torch.cuda.synchronize()
    ts.append(e0.elapsed_time(e1))
ms = sorted(ts)[len(ts) // 2]
flop = 151_912_448 * R
print(json.dumps({"rays": R, "precision": prec, "ms": ms, "all_ms": ts, "tflops": flop / ms / 1e9, "rays_per_s": R / ms * 1e3}))
