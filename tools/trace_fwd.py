"""Debug aid: clock64 stamps of CTA 0 of the fused forward.  Epilogue stamps of the first ray pair (per layer: accumulators
ready / row handed back, tile 0) and the issuer's accumulated wait cycles over the whole launch."""
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
params = [sd[k].to(dev) for k in _cabi.PARAM_ORDER]
d = torch.nn.functional.normalize(torch.randn(R, 3) + torch.tensor([0., 0., -2.]), dim=-1).to(dev)
o = torch.zeros(3, device=dev)
packed = _ops.pack_weights(params)
rc = _ops.make_cfg(R, 128, 0.01, 10.0, 1, 0, 0, 1, _cabi.PREC_BF16)
tr = torch.zeros(512, dtype=torch.int64, device=dev)
save = "--save" in sys.argv    # trace the saving (training) forward instead of the render forward
if save:
    params = [p.requires_grad_(True) for p in params]
    noise = torch.rand(R, 128, device=dev)
    run = lambda: _ops.RenderRays.apply(rc, packed, o, d, noise, *params)
else:
    run = lambda: _ops.render_nograd(rc, params, packed, o, d, None, want_aux=False)
for _ in range(2):
    run()
_cabi.call("nnb_debug_set_trace", _cabi.ptr(tr))
run()
torch.cuda.synchronize()
_cabi.call("nnb_debug_set_trace", None)
full = tr.cpu()
epi = full[:64].reshape(16, 4)
t0 = int(epi[0, 0])
print("layer: acc_ready  row_handed_back   (tile 0, first pair; cycles from layer 0's acc_ready)")
for L in range(9):
    print("%2d: %8d %8d" % (L, int(epi[L, 0]) - t0, int(epi[L, 1]) - t0))
w = [int(x) for x in full[100:105]]
if w[3]:
    print("issuer (CTA 0, whole launch, %d ray pairs): total %d cycles; waiting for activations %d (%.0f%%), own weights %d (%.0f%%), "
          "peer weights %d (%.0f%%)" % (w[4], w[3], w[0], 100.0 * w[0] / w[3], w[1], 100.0 * w[1] / w[3], w[2], 100.0 * w[2] / w[3]))
