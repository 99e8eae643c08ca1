"""Times nnb_chamfer_nn in isolation on the very clouds the cfg2 training step feeds it (captured from one eager step)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "my-nope-nerf_b200")):
    sys.path.insert(0, p)
import torch
import bench, synthetic_scene as syn
from model import parallel, _ops
dev = torch.device("cuda:0")
cfg, scene, tr = bench.build_trainer(dev, 20, 4096, "bf16", parallel.DataParallelContext(0, 1))
captured = []
orig = _ops.chamfer_nn


def spy(src, dst, want_dist=False):
    captured.append((src.detach().clone(), dst.detach().clone()))
    return orig(src, dst, want_dist=want_dist)


_ops.chamfer_nn = spy
b = syn.make_batch(scene, 3, 4)
b = {k: (v.to(dev) if torch.is_tensor(v) and k not in ("img.idx", "img.ref_idxs") else v) for k, v in b.items()}
tr.train_step(b, it=0, epoch=0, scheduling_start=cfg["training"]["scheduling_start"])
_ops.chamfer_nn = orig
torch.cuda.synchronize()
for i, (s, d) in enumerate(captured):
    for _ in range(3):
        orig(s, d)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        idx, dist = orig(s, d, want_dist=True)
    e1.record()
    torch.cuda.synchronize()
    print("direction %d: %d x %d points, %.3f ms per call; NN distance median %.4f, p99 %.4f, max %.4f; extent %.2f"
          % (i, s.shape[0], d.shape[0], e0.elapsed_time(e1) / 20, float(dist.median()), float(dist.quantile(0.99)), float(dist.max()),
             float((d - d[0]).abs().max())))
