"""Prints per-layer clock64 stamps of CTA 0 / first ray pair of the fused forward (debug aid)."""
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
tr = torch.zeros(256, dtype=torch.int64, device=dev)
for _ in range(2):
    _ops.render_nograd(rc, params, packed, o, d, None, want_aux=False)
_cabi.call("nnb_debug_set_trace", _cabi.ptr(tr))
_ops.render_nograd(rc, params, packed, o, d, None, want_aux=False)
torch.cuda.synchronize()
_cabi.call("nnb_debug_set_trace", None)
full = tr.cpu()
epi = full[:64].reshape(16, 4)
mma = full[64:]
starts = [0, 3, 12, 21, 30, 41, 50, 59, 68, 77, 86]   # first issue group of each layer (csrc/mlp_tc.cu build_groups)
t0 = int(mma[0])
print("layer: act_ready(mma)  last_group_issued(mma)  acc_ready(epi)  epi_done   | mma_span  commit->epi  epi_span  epi->next_mma")
for L in range(10):
    a, b = int(mma[2 * starts[L]]) - t0, int(mma[2 * (starts[L + 1] - 1) + 1]) - t0
    c, e = int(epi[L, 0]) - t0, int(epi[L, 1]) - t0
    nxt = int(mma[2 * starts[L + 1]]) - t0 if L < 9 else 0
    print("%2d: %8d %8d %8d %8d | %6d %6d %6d %6d" % (L, a, b, c, e, b - a, c - b, e - c, nxt - e))
