"""Debug aid: per-CTA globaltimer stamps of the weight-gradient kernel (start, streaming done, flushed), grouped by job pair."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "my-nope-nerf_b200")):
    sys.path.insert(0, p)
import torch
import synthetic_scene as syn
from model import _cabi, _ops
R = 4096
dev = torch.device("cuda:0")
sd = syn.init_mlp_state(42)
params = [sd[k].to(dev).requires_grad_(True) for k in _cabi.PARAM_ORDER]
d = torch.nn.functional.normalize(torch.randn(R, 3) + torch.tensor([0., 0., -2.]), dim=-1).to(dev)
o = torch.zeros(3, device=dev)
packed = _ops.pack_weights(params)
rc = _ops.make_cfg(R, 128, 0.01, 10.0, 1, 0, 0, 1, _cabi.PREC_BF16)
noise = torch.rand(R, 128, device=dev)
tr = torch.zeros(4096, dtype=torch.int64, device=dev)
def run():
    out = _ops.RenderRays.apply(rc, packed, o, d, noise, *params)
    (out[0].sum() + out[1].sum()).backward()
for _ in range(2):
    run()
_cabi.call("nnb_debug_set_trace", _cabi.ptr(tr))
run()
torch.cuda.synchronize()
_cabi.call("nnb_debug_set_trace", None)
t = tr.cpu()[1024:1024 + 4 * 296].reshape(296, 4)
t = t[t[:, 0] > 0]
t0 = int(t[:, 0].min())
print("kernel span %.1f us (first start -> last flush)" % ((int(t[:, 2].max()) - t0) / 1e3))
print("job pair: CTAs | start min..max | streaming done min..max | flushed min..max  (us)")
for p in sorted(set(int(x) for x in t[:, 3])):
    m = t[:, 3] == p
    a, b, c = [(t[m, k] - t0).double() / 1e3 for k in range(3)]
    print("%2d: %3d | %6.1f..%6.1f | %7.1f..%7.1f | %7.1f..%7.1f" % (p, int(m.sum()), a.min(), a.max(), b.min(), b.max(), c.min(), c.max()))
