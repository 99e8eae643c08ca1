"""Where does the end-to-end leg of bench.py lose time?  (1) H2D bandwidth of one batch from pinned memory, alone;
(2) the same upload while a training step runs on the compute stream."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "my-nope-nerf_b200")):
    sys.path.insert(0, p)
import torch
import bench, synthetic_scene as syn
from model import parallel
import model as mdl
dev = torch.device("cuda:0")
dp = parallel.DataParallelContext(0, 1)
cfg, scene, tr = bench.build_trainer(dev, 20, 4096, "bf16", dp)
n = scene["n"]
sched = cfg["training"]["scheduling_start"]
host = []
for i in range(n):
    b = syn.make_batch(scene, i, i + 1 if i < n - 1 else i - 1)
    host.append({k: (v.pin_memory() if torch.is_tensor(v) else v) for k, v in b.items()})
nbytes = bench.batch_bytes(host[0])
rs = mdl.ResidentScene(host, dev)
devb = [rs.batch(i, i + 1 if i < n - 1 else i - 1) for i in range(n)]
for it in list(range(5)) + [n - 2, n - 1]:
    tr.train_step_graphed(devb[it % n], it=it, epoch=0, scheduling_start=sched)
torch.cuda.synchronize()
cs = torch.cuda.Stream()


def upload(b):
    with torch.cuda.stream(cs):
        return {k: (v.to(dev, non_blocking=True) if torch.is_tensor(v) and k not in ("img.idx", "img.ref_idxs") else v) for k, v in b.items()}


def ev():
    return torch.cuda.Event(enable_timing=True)


for label, busy in (("alone", False), ("while a step runs", True)):
    ts = []
    for r in range(6):
        torch.cuda.synchronize()
        if busy:
            tr.train_step_graphed(devb[r % 4], it=r, epoch=0, scheduling_start=sched)
        e0, e1 = ev(), ev()
        e0.record(cs)
        d = upload(host[r % n])
        e1.record(cs)
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts = sorted(ts)[1:-1]
    ms = sum(ts) / len(ts)
    print("H2D of one batch (%d tensors, %.1f MB) %s: %.3f ms = %.1f GB/s" % (sum(torch.is_tensor(v) for v in host[0].values()), nbytes / 1e6, label, ms, nbytes / ms / 1e6))
for k, v in host[0].items():
    if torch.is_tensor(v):
        print("   %-18s %10d B pinned=%s" % (k, v.numel() * v.element_size(), v.is_pinned()))
