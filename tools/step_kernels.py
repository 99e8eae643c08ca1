"""Chronological list of the GPU activities of one graphed training step (name, start offset, duration)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "my-nope-nerf_b200")):
    sys.path.insert(0, p)
import torch
import bench, synthetic_scene as syn
from model import parallel
dev = torch.device("cuda:0")
dp = parallel.DataParallelContext(0, 1)
cfg, scene, tr = bench.build_trainer(dev, 20, 4096, "bf16", dp)
sched = cfg["training"]["scheduling_start"]
batches = []
for i in range(4):
    b = syn.make_batch(scene, i, i + 1)
    batches.append({k: (v.to(dev) if torch.is_tensor(v) and k not in ("img.idx", "img.ref_idxs") else v) for k, v in b.items()})
step = tr.train_step if '--no-graph' in sys.argv else tr.train_step_graphed
for it in range(6):
    step(batches[it % 4], it=it, epoch=0, scheduling_start=sched)
torch.cuda.synchronize()
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    step(batches[2], it=6, epoch=0, scheduling_start=sched)
    torch.cuda.synchronize()
evs = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
evs.sort(key=lambda e: e.time_range.start)
t0 = evs[0].time_range.start
busy = 0.0
for e in evs:
    d = e.time_range.end - e.time_range.start
    busy += d
    print("%9.1f %8.1f s%-3s %s" % (e.time_range.start - t0, d, getattr(e, "device_resource_id", "?"), e.name[:100]))
print("span %.1f us, busy %.1f us, %d activities" % (evs[-1].time_range.end - t0, busy, len(evs)))
