"""Where does a training step's wall time go?  CPU launch time vs GPU busy time, via torch.profiler."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "my-nope-nerf_b200")):
    sys.path.insert(0, p)
import torch
import bench, synthetic_scene as syn
from model import parallel
dev = torch.device("cuda:0")
dp = parallel.DataParallelContext(0, 1)
if '--fork-config' in sys.argv:
    # the fork's own settings (bench_extras.fork_config): 188 x 621, 1,024 rays, sparse depth, frozen poses, host batches, torch Adam
    import bench_extras as X
    cfg = syn.default_cfg()
    scene, batches = X.fork_scene(20)
    batches = batches[:4]
    tr = X.fork_trainer(dev, cfg, scene, 1024)
else:
    cfg, scene, tr = bench.build_trainer(dev, 20, 4096, "bf16", dp)
    batches = []
    for i in range(4):
        b = syn.make_batch(scene, i, i + 1)
        batches.append({k: (v.to(dev) if torch.is_tensor(v) and k not in ("img.idx", "img.ref_idxs") else v) for k, v in b.items()})
n = scene["n"]
sched = cfg["training"]["scheduling_start"]
if '--no-graph' in sys.argv:
    tr.use_graph = False          # kernel by kernel (what a batch that does not qualify for the graph gets)
step = tr.train_step
for it in range(5):
    step(batches[it % 4], it=it, epoch=0, scheduling_start=sched)
torch.cuda.synchronize()
import time
t0 = time.perf_counter()
for it in range(20):
    step(batches[it % 4], it=it, epoch=0, scheduling_start=sched)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print("cpu issue time per step %.3f ms; wall per step incl. drain %.3f ms" % ((t1 - t0) / 20 * 1e3, (t2 - t0) / 20 * 1e3))
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for it in range(3):
        step(batches[it % 4], it=it, epoch=0, scheduling_start=sched)
    torch.cuda.synchronize()
ev = prof.key_averages()
rows = sorted(ev, key=lambda e: -e.self_cpu_time_total)[:25]
for e in rows:
    print("%-60s cpu %8.1f us  n=%d" % (e.key[:60], e.self_cpu_time_total / 3, e.count // 3))
tot_cuda = sum(e.self_device_time_total for e in ev) / 3
print("GPU kernel time per step: %.3f ms" % (tot_cuda / 1e3))
rows = sorted(ev, key=lambda e: -e.self_device_time_total)[:30]
print("--- GPU time per step by kernel")
for e in rows:
    if e.self_device_time_total > 0:
        print("%-90s gpu %8.1f us  n=%d" % (e.key[:90], e.self_device_time_total / 3, e.count // 3))
if '--trace' in sys.argv:
    # chronological device activity of the LAST profiled step: start offset, duration, stream, name; plus idle gaps on the union
    path = os.path.join(ROOT, "gpurun_out", "step_trace.json")
    os.makedirs(os.path.dirname(path), exist_ok=True)
    prof.export_chrome_trace(path)
    tr_ev = [e for e in json.load(open(path))["traceEvents"] if e.get("cat") in ("kernel", "gpu_memcpy", "gpu_memset")]
    tr_ev.sort(key=lambda e: e["ts"])
    n3 = len(tr_ev) // 3
    last = tr_ev[2 * n3:]
    t0 = last[0]["ts"]
    print("--- last step, chronological (us from its first device activity): start dur stream name")
    busy_end = t0
    idle = 0.0
    for e in last:
        if e["ts"] > busy_end:
            idle += e["ts"] - busy_end
        gap = e["ts"] - busy_end
        busy_end = max(busy_end, e["ts"] + e["dur"])
        print("%9.1f %8.1f s%-3s %s%s" % (e["ts"] - t0, e["dur"], e["args"].get("stream", "?"), e["name"][:80],
                                         ("   <- idle %.1f before" % gap) if gap > 1.5 else ""))
    print("span %.1f us, device idle (no stream busy) %.1f us" % (busy_end - t0, idle))
    os.remove(path)
