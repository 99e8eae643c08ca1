"""Times nnb_chamfer_nn (both screening passes) on depth-cloud shaped inputs.  usage: python tools/time_chamfer.py [P]"""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "my-nope-nerf_b200"))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from model import _ops, _cabi as K

P = int(sys.argv[1]) if len(sys.argv) > 1 else 28830
dev = torch.device("cuda:0")
g = torch.Generator().manual_seed(0)
a = torch.rand(P, 3, generator=g)
a = torch.stack([(a[:, 0] * 2 - 1) * 2, (a[:, 1] * 1.5 - 0.75) * 2, -(1.5 + a[:, 2])], 1).to(dev)
b = (a + 0.01 * torch.randn(P, 3, generator=g).to(dev))[torch.randperm(P, generator=g).to(dev)]
if "--raster" in sys.argv:
    # back-projected depth maps in raster order (what the training step feeds): two smooth surfaces a small rigid offset apart
    hh = int(round((P / 1.2) ** 0.5))
    ww = (P + hh - 1) // hh
    v, u = torch.meshgrid(torch.linspace(-0.75, 0.75, hh), torch.linspace(-1, 1, ww), indexing="ij")
    zz = 2.0 + 0.3 * torch.sin(3 * u) * torch.cos(2 * v)
    a = torch.stack([u * zz, v * zz, -zz], -1).reshape(-1, 3)[:P]
    b = (a + torch.tensor([0.013, -0.007, 0.02]) + 1e-3 * torch.randn(a.shape, generator=g)).to(dev)
    a = a.to(dev)
for mode, name in ((0, "tcgen05 both"), (2, "tcgen05 + cc exact"), (1, "cuda cores both")):
    K.lib().nnb_debug_chamfer_pass1(mode)
    for _ in range(3):
        idx, d = _ops.chamfer_nn(a, b, want_dist=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n = 20
    e0.record()
    for _ in range(n):
        idx, d = _ops.chamfer_nn(a, b, want_dist=True)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / n
    print("%-18s P=%d: %.3f ms per direction, %.2f Gpairs/s, checksum %d" % (name, P, ms, P * P / ms / 1e6, int(idx.sum())))
K.lib().nnb_debug_chamfer_pass1(0)
