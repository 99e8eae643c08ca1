"""Extra legs of bench.py (everything in `extras` and the per-stage rooflines): device-timed with CUDA events, medians of
several repeats, each on >= 65,536 units where the stage allows it so that the figure is not launch latency.

Nothing here is on the product path.  `torch_eager_gpu` executes oracle/ (the plain-PyTorch restatement of the reference) as a
COMPARATOR on the same GPU -- SURVEY.md section 0 names stock PyTorch eager on the same device "the comparison target"."""
import time

import torch

import synthetic_scene as syn

FLOP_PER_RAY_TRAIN = 455_737_344
FLOP_PER_RAY_RENDER = 151_912_448


def med_ms(fn, reps=7, warm=2):
    """Median device time of fn() in ms (CUDA events on the current stream, synchronised on both sides)."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]


def graph_ms(fn, inner=10, reps=7):
    """Median DEVICE time of one fn() in ms: `inner` calls captured in a CUDA graph and replayed, so that no host launch
    overhead sits between the events (a 10 us kernel launched from Python is otherwise timed at 30-60 us)."""
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for _ in range(2):
            fn()
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(inner):
            fn()
    g.replay()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) / inner)
    del g
    return sorted(ts)[len(ts) // 2]


def _fresh_trainer(dev, scene, cfg, rays_global, precision, dp, flat=True, **tkw):
    import model as mdl
    n = scene["n"]
    nerf = mdl.OfficialStaticNerf(cfg)
    nerf.load_state_dict(syn.init_mlp_state(42))
    net = mdl.get_model(mdl.Renderer(nerf, dict(cfg["rendering"], precision=precision), device=dev), cfg, device=dev).to(dev)
    pose = mdl.LearnPose(n, True, True, cfg, init_c2w=scene["poses_gt"].clone()).to(dev)
    dist = mdl.Learn_Distortion(n, True, True, cfg).to(dev)
    r0, t0 = syn.init_pose_params(n)
    with torch.no_grad():
        pose.r.copy_(r0)
        pose.t.copy_(t0)
    tcfg = dict(cfg["training"], n_training_points=rays_global, **tkw)
    Adam = mdl.FlatAdam if flat else torch.optim.Adam
    tr = mdl.Trainer(net, Adam(net.parameters(), lr=tcfg["learning_rate"]), tcfg, device=dev,
                     optimizer_pose=Adam(pose.parameters(), lr=tcfg["pose_lr"]), pose_param_net=pose,
                     optimizer_distortion=Adam(dist.parameters(), lr=tcfg["distortion_lr"]), distortion_net=dist, dp=dp)
    return tr, nerf, pose, dist


def release(tr):
    """Drop a trainer's captured graphs (each owns a private pool with a saved-activation workspace) and return the memory."""
    if hasattr(tr, "_graphs"):
        tr._graphs.clear()
    import gc
    gc.collect()
    torch.cuda.empty_cache()


# ---------------------------------------------------------------------------------------------------------------------
# the path an UNCHANGED train.py gets (train.py:204-219): DataLoader-shaped pinned host batches handed to Trainer.train_step,
# three torch.optim.Adam, and the three .item() reads of every step
# ---------------------------------------------------------------------------------------------------------------------
def dropin_path(dev, scene, cfg, host_batches, precision, dp, steps=40):
    from model import parallel
    tr, *_ = _fresh_trainer(dev, scene, cfg, 4096 * dp.world_size, precision, dp, flat=False)
    n = scene["n"]
    sched = cfg["training"]["scheduling_start"]
    for it in list(range(n)) + [0, 1]:          # every graph variant captured, allocator settled
        tr.train_step(host_batches[it % n], it, 0, sched, None)
    torch.cuda.synchronize()
    if dp.world_size > 1:
        torch.distributed.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    sink = 0.0
    for it in range(steps):
        loss_dict = tr.train_step(host_batches[it % n], it, 0, sched, None)
        sink += loss_dict['l2_mean'].item() + loss_dict['loss_pc'].item() + loss_dict['loss_rgb_s'].item()
    e1.record()
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) * 1e3 / steps
    ms = e0.elapsed_time(e1) / steps
    out = {"ms_per_step": ms, "wall_ms_per_step": wall, "rays_per_s": 4096 * dp.world_size / (ms * 1e-3), "steps": steps,
           "what": "Trainer.train_step(batch, it, epoch, scheduling_start, render_path) exactly as train.py:213 calls it: pinned HOST "
                   "batches, 3 x torch.optim.Adam, .item() on l2_mean / loss_pc / loss_rgb_s every step (train.py:216-218), no "
                   "precision key in the YAML (-> bf16), the step replayed as a CUDA graph by train_step itself",
           "graph_variants": len(getattr(tr, "_graphs", {})), "finite": bool(sink == sink)}
    release(tr)
    return out


# ---------------------------------------------------------------------------------------------------------------------
# stock PyTorch eager on the same GPU: the oracle port (plain torch ops, the reference's algorithm) on cuda
# ---------------------------------------------------------------------------------------------------------------------
def torch_eager_gpu(dev, scene, cfg, steps=3):
    from oracle import nope_oracle as O
    n = scene["n"]
    sd = syn.init_mlp_state(42)
    P = {k: v.clone().to(dev).requires_grad_(True) for k, v in sd.items()}
    r0, t0 = syn.init_pose_params(n)
    pose = {"r": r0.clone().to(dev).requires_grad_(True), "t": t0.clone().to(dev).requires_grad_(True), "init_c2w": scene["poses_gt"].to(dev)}
    dist = {"global_scales": torch.ones(n, 1, device=dev, requires_grad=True), "global_shifts": torch.zeros(n, 1, device=dev, requires_grad=True)}
    params = list(P.values()) + [pose["r"], pose["t"], dist["global_scales"], dist["global_shifts"]]
    opt = torch.optim.Adam(params, lr=1e-3)
    batch = {k: (v.to(dev) if torch.is_tensor(v) else v) for k, v in syn.make_batch(scene, 0, 1).items()}
    R = 4096
    ray_idx = syn.ray_choice(scene["h"], scene["w"], R).to(dev)
    noise = syn.jitter(R).to(dev)
    ts = []
    torch.cuda.reset_peak_memory_stats()
    with torch.device(dev):          # the oracle's tensor factories follow the default device
        for it in range(steps + 1):
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            opt.zero_grad()
            ld, _, _ = O.train_loss(batch, P, pose, dist, cfg, ray_idx, noise, epoch=0)
            ld["loss"].backward()
            opt.step()
            e1.record()
            torch.cuda.synchronize()
            if it > 0:
                ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    out = {"ms_per_step": ms, "rays_per_s": R / (ms * 1e-3), "steps": steps, "loss": float(ld["loss"]),
           "peak_memory_gb": torch.cuda.max_memory_allocated() / 1e9,
           "what": "one full cfg2 step (4,096 rays x 128 samples, full loss set, backward, one torch.optim.Adam over all parameters) of "
                   "the oracle port = the reference's algorithm as plain PyTorch fp32 ops, eager, on this GPU (TF32 off); the "
                   "Chamfer argmin is evaluated on source slices like the reference's own split"}
    del P, pose, dist, opt, ld
    torch.cuda.empty_cache()
    return out


# ---------------------------------------------------------------------------------------------------------------------
# cfg3: the 65,536-ray GLOBAL batch sharded over the N ranks (strong scaling: 65,536 / N rays per GPU)
# ---------------------------------------------------------------------------------------------------------------------
def fork_scene(n=20, h=188, w=621):
    scene = syn.make_scene("cfg2", frames=n, h=h, w=w, sparse_depth=True)
    host = []
    for i in range(n):
        b = syn.make_batch(scene, i, i + 1 if i < n - 1 else i - 1, dpt=False)
        host.append({k: (v.pin_memory() if torch.is_tensor(v) else v) for k, v in b.items()})
    return scene, host


def fork_trainer(dev, cfg, scene, rays):
    import model as mdl
    n = scene["n"]
    nerf = mdl.OfficialStaticNerf(cfg)
    nerf.load_state_dict(syn.init_mlp_state(42))
    net = mdl.get_model(mdl.Renderer(nerf, cfg["rendering"], device=dev), cfg, device=dev).to(dev)
    pose = mdl.LearnPose(n, False, False, cfg, init_c2w=scene["poses_gt"].clone()).to(dev)
    dist = mdl.Learn_Distortion(n, False, False, cfg).to(dev)
    tcfg = dict(cfg["training"], n_training_points=rays)
    return mdl.Trainer(net, torch.optim.Adam(net.parameters(), lr=tcfg["learning_rate"]), tcfg, device=dev,
                       optimizer_pose=torch.optim.Adam(pose.parameters(), lr=tcfg["pose_lr"]), pose_param_net=pose,
                       optimizer_distortion=torch.optim.Adam(dist.parameters(), lr=tcfg["distortion_lr"]), distortion_net=dist)


def fork_config(dev, cfg, steps=60):
    """The fork's OWN training configuration (configs/V_KITTI/*.yaml over configs/default.yaml) rather than BASELINE's cfg2: frames at
    half resolution (`resize_factor: 2` -> 188 x 621), 1,024 rays per step (`n_training_points` default), sparse ground-truth depth
    (`with_depth`, `sparsify_depth_pattern [1,9,1,9]` -> `img.depth` + validity mask), poses and distortion fixed at their initial
    values (`learn_R` / `learn_t` / `learn_scale` / `learn_shift: false`), three torch.optim.Adam, pinned host batches through
    `Trainer.train_step` with the three `.item()` reads of train.py:216-218.  Timed twice: as `train_step` routes it (CUDA graph;
    sparse-depth batches qualify since the in-graph re-draw) and kernel by kernel (`use_graph = False`)."""
    n, R = 20, 1024
    scene, host = fork_scene(n)
    sched = cfg["training"]["scheduling_start"]
    out = {"what": "188 x 621 frames, 1,024 rays x 128 samples, sparse img.depth (1 pixel in 100 valid), poses / distortion frozen, full loss "
                   "set, 3 x torch.optim.Adam, host batches through Trainer.train_step + 3 .item() per step"}
    for mode in ("graph", "kernel_by_kernel"):
        tr = fork_trainer(dev, cfg, scene, R)
        if mode == "kernel_by_kernel":
            tr.use_graph = False
        for it in list(range(n)) + [0, 1]:
            tr.train_step(host[it % n], it, 0, sched, None)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        sink = 0.0
        for it in range(steps):
            ld = tr.train_step(host[it % n], it, 0, sched, None)
            sink += ld['l2_mean'].item() + ld['loss_pc'].item() + ld['loss_rgb_s'].item()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        out[mode] = {"ms_per_step": ms, "rays_per_s": R / (ms * 1e-3), "finite": bool(sink == sink),
                     "graph_variants": len(getattr(tr, "_graphs", {}))}
        release(tr)
    out["speedup_of_the_graph"] = out["kernel_by_kernel"]["ms_per_step"] / out["graph"]["ms_per_step"]
    return out


def cfg3(dev, scene, cfg, host_batches, resident, precision, dp, steps=10, repeats=3):
    """One graph variant only (frames 0..n-3): at N = 1 the saved-activation workspace of a 65,536-ray step is 86 GB and every
    captured variant owns one."""
    G = 65536
    world = dp.world_size
    tr, *_ = _fresh_trainer(dev, scene, cfg, G, precision, dp)
    n = scene["n"]
    frames = list(range(n - 2))
    sched = cfg["training"]["scheduling_start"]
    # same keys AND shapes as the end-to-end leg's uploaded batches, so that both legs replay ONE captured graph (a second variant
    # would own a second 86 GB workspace at N = 1)
    keys = [k for k in host_batches[0] if k in tr.STEP_KEYS]
    dev_batches = [{k: (host_batches[i][k].to(dev) if torch.is_tensor(host_batches[i][k]) and k not in ("img.idx", "img.ref_idxs") else host_batches[i][k])
                    for k in keys} for i in frames]

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            torch.distributed.barrier()
            torch.cuda.synchronize()

    for it in range(4):
        tr.train_step_graphed(dev_batches[it % len(frames)], it=it, epoch=0, scheduling_start=sched)
    dp.timing = []
    ms_all = []
    for _ in range(repeats):
        sync_all()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for it in range(steps):
            tr.train_step_graphed(dev_batches[it % len(frames)], it=it, epoch=0, scheduling_start=sched)
        e1.record()
        sync_all()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms_all.append(float(t) / steps)
    ar = [a.elapsed_time(b) for a, b in dp.timing]
    dp.timing = None
    # end to end: the step's host batch (pinned) copied to the device inside the timed region, loss read back every step
    from model.scene import BatchUploader
    up = BatchUploader(tr, host_batches[0])
    loss_host = torch.zeros(1).pin_memory()
    sync_all()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    up.submit(host_batches[frames[0]])
    for it in range(steps):
        cur = up.next()
        if it + 1 < steps:
            up.submit(host_batches[frames[(it + 1) % len(frames)]])
        ld = tr.train_step_graphed(cur, it=it, epoch=0, scheduling_start=sched)
        up.done()
        loss_host.copy_(ld["loss"].detach().reshape(1), non_blocking=True)
    e1.record()
    sync_all()
    t = torch.tensor([e0.elapsed_time(e1)], device=dev)
    if world > 1:
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    ms = sorted(ms_all)[len(ms_all) // 2]
    ms_e2e = float(t) / steps
    assert len(tr._graphs) == 1, "cfg3 legs must share one graph variant"
    out = {"global_rays": G, "rays_per_gpu": G // world, "n_gpus": world, "ms_per_step": ms, "ms_per_step_repeats": [round(m, 3) for m in ms_all],
           "rays_per_s": G / (ms * 1e-3), "e2e": {"ms_per_step": ms_e2e, "rays_per_s": G / (ms_e2e * 1e-3)},
           "tflops_per_gpu": FLOP_PER_RAY_TRAIN * (G / world) / (ms * 1e-3) / 1e12, "scaling": "strong", "steps": steps,
           "workspace_gb_per_gpu": 1.31e6 * (G / world) / 1e9}
    if ar:
        ar = sorted(ar)
        out["allreduce_in_step_us"] = 1e3 * ar[len(ar) // 2]
    release(tr)
    return out


# ---------------------------------------------------------------------------------------------------------------------
# the gradient allreduce alone (SURVEY.md 8d last row): one flat fp32 buffer of MLP + pose + distortion gradients
# ---------------------------------------------------------------------------------------------------------------------
def allreduce_alone(dev, dp, n_frames, reps=30):
    if dp.world_size == 1:
        return None
    numel = 595_844 + 8 * n_frames + 10          # + the loss scalars that ride along
    buf = torch.randn(numel, device=dev)
    dist = torch.distributed
    for _ in range(5):
        dist.all_reduce(buf, op=dist.ReduceOp.AVG)
    ts = []
    for _ in range(reps):
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        dist.all_reduce(buf, op=dist.ReduceOp.AVG)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ts.append(float(t))
    ms = sorted(ts)[len(ts) // 2]
    nbytes = numel * 4
    W = dp.world_size
    return {"bytes": nbytes, "us": ms * 1e3, "bus_gb_per_s": 2.0 * (W - 1) / W * nbytes / (ms * 1e-3) / 1e9,
            "nominal_gb_per_s_per_direction": 900.0,
            "note": "one NCCL all_reduce(AVG) of the flat fp32 gradient buffer, CUDA events, max over ranks, median of %d; at 2.4 MB "
                    "the collective is latency-bound, not bandwidth-bound" % reps}


# ---------------------------------------------------------------------------------------------------------------------
# data-parallel parity over NCCL: the all-reduced gradient of the sharded step vs the one-rank gradient of the SAME global batch
# ---------------------------------------------------------------------------------------------------------------------
def dp_parity(dev, scene, cfg, dp, rays_global=2048):
    """Every rank runs compute_loss + backward on its shard and the gradients are all-reduced (the product path).  Rank 0 then
    repeats the step alone on the whole batch (untimed) and reports the largest scale-relative difference per tensor, for both
    precision paths and both depth-loss types."""
    from model import parallel
    if dp.world_size == 1:
        return None
    h, w, n = scene["h"], scene["w"], scene["n"]
    ray_idx = syn.ray_choice(h, w, rays_global).to(dev)
    noise = syn.jitter(rays_global).to(dev)
    batch = {k: (v.to(dev) if torch.is_tensor(v) and k not in ("img.idx", "img.ref_idxs") else v) for k, v in syn.make_batch(scene, 3, 4).items()}
    sched = cfg["training"]["scheduling_start"]
    out = {"global_rays": rays_global, "n_gpus": dp.world_size}

    def step(dpx, precision, dlt):
        tr, nerf, pose, dist = _fresh_trainer(dev, scene, cfg, rays_global, precision, dpx, flat=False, depth_loss_type=dlt)
        tr.fixed_ray_idx, tr.fixed_noise = ray_idx, noise
        ld = tr.compute_loss(batch, it=1, epoch=0, scheduling_start=sched)
        ld["loss"].backward()
        params = [p for m in (nerf, pose, dist) for p in m.parameters()]
        scal = tr._loss_scalars(ld)
        if dpx.world_size > 1:
            dpx.allreduce_grads(params, extra=scal)
        torch.cuda.synchronize()
        return [p.grad.detach().clone() if p.grad is not None else torch.zeros_like(p) for p in params], {k: float(ld[k]) for k in ("loss", "loss_rgb", "loss_depth", "loss_pc", "loss_rgb_s")}

    for precision in ("fp32", "bf16"):
        for dlt in ("l1", "invariant"):
            g_dp, l_dp = step(dp, precision, dlt)
            if dp.rank == 0:
                g_1, l_1 = step(parallel.DataParallelContext(0, 1), precision, dlt)
                worst = max(float((a - b).abs().max() / b.abs().max().clamp_min(1e-20)) for a, b in zip(g_dp, g_1))
                num = sum(float(((a - b).double() ** 2).sum()) for a, b in zip(g_dp, g_1))
                den = sum(float((b.double() ** 2).sum()) for b in g_1)
                out["%s_%s" % (precision, dlt)] = {
                    "grad_max_scale_relative_diff": worst, "grad_rel_l2": (num / max(den, 1e-300)) ** 0.5,
                    "loss_rel_diff": max(abs(l_dp[k] - l_1[k]) / max(abs(l_1[k]), 1e-12) for k in l_1)}
            torch.distributed.barrier()
    out["note"] = ("gradients after the NCCL allreduce of the sharded step vs rank 0 recomputing the whole batch alone; fp32: only the order "
                   "of fp32 additions differs; bf16: per-ray results are identical on every rank, the weight-gradient atomics add in another order")
    return out


# ---------------------------------------------------------------------------------------------------------------------
# per-stage rooflines of the HBM-bound stages: algorithmic bytes per unit x units / CUDA-event time
# ---------------------------------------------------------------------------------------------------------------------
def stage_rooflines(dev, hbm_gbs, units=4194304):
    """Each stage on `units` (>= 65,536; default 4 M rays, 2 M points) rays / points, timed as CUDA-graph replays (graph_ms: device time, no
    host launch gaps), so that the figure is bandwidth and not launch latency.  Bytes per unit are SURVEY.md 8d's algorithmic
    figures (inputs read once + outputs written once).  Between replays the inputs + outputs of a stage (tens to hundreds of
    MB) exceed or rival the 126 MB L2 only for the compositing stages; the smaller ones are L2-resident on replay, which the
    `note` of each entry says -- their figure is an upper bound of what HBM would give."""
    from model import _ops
    g = torch.Generator().manual_seed(5)
    R = units
    h, w = 2048, 4096                                   # source image of the gathers: 8 M pixels
    img = torch.rand(3, h, w, generator=g).to(dev)
    depth_img = (torch.rand(h, w, generator=g) + 0.5).to(dev)
    ray_idx = torch.randint(0, h * w, (R,), generator=g).to(dev)
    ss = torch.tensor([1.1, 0.02], device=dev)
    K = torch.diag(torch.tensor([1.17, -3.87, -1.0, 1.0])).unsqueeze(0).to(dev)
    c2w = syn.make_poses(3)[2].to(dev)
    W = torch.inverse(c2w).unsqueeze(0).contiguous()
    S = torch.eye(4, device=dev).unsqueeze(0)
    pixels, rgb_gt, depth_sel = _ops.GatherRays.apply(ray_idx, img, depth_img, ss, h, w)
    out = []
    med_ms = graph_ms      # noqa: F811  (every stage below is timed as a graph replay)

    def add(name, ms, bytes_per_unit, n_units, unit, note=""):
        gbs = bytes_per_unit * n_units / (ms * 1e-3) / 1e9
        mb = bytes_per_unit * n_units / 1e6
        out.append({"kernel": name, "bound": "hbm", "ms": ms, "us": ms * 1e3, "units": n_units, "unit": "GB/s", "per": unit,
                    "bytes_per_unit": bytes_per_unit, "achieved": gbs, "peak": hbm_gbs, "frac": gbs / hbm_gbs,
                    "note": note + ("; %.0f MB per launch%s" % (mb, " (fits the 126 MB L2 on replay: an upper bound of what HBM gives)" if mb < 126 else ""))})

    with torch.no_grad():
        add("nnb_gather_rays_fwd (ray gather: pixels, rgb_gt, depth*scale+shift)", med_ms(lambda: _ops.GatherRays.apply(ray_idx, img, depth_img, ss, h, w)),
            8 + 12 + 4 + 8 + 12 + 4, R, "ray", "read idx 8 + rgb 12 + depth 4 (scattered 4-byte reads: 32-byte sectors are the real traffic), write pixels 8 + rgb 12 + depth 4")
        add("nnb_raygen_fwd (stage 1)", med_ms(lambda: _ops.RayGen.apply(pixels.unsqueeze(0), depth_sel.reshape(1, -1, 1), K, W, S, True)),
            8 + 4 + 12 + 4 + 4 + 1, R, "ray", "SURVEY 8d: 33 B/ray")
        pred = torch.rand(1, R, 3, generator=g).to(dev)
        add("nnb_rgb_loss_fwd", med_ms(lambda: _ops.RgbLoss.apply(pred, rgb_gt.unsqueeze(0))), 24, R, "ray", "SURVEY 8d: 24 B/ray")
        dp_, dg_ = (torch.rand(R, generator=g) + 0.5).to(dev), (torch.rand(R, generator=g) + 0.5).to(dev)
        mask = torch.ones(R, dtype=torch.bool, device=dev)
        add("nnb_depth_l1_fwd", med_ms(lambda: _ops.DepthL1.apply(dp_, dg_, mask)), 9, R, "ray", "SURVEY 8d: 8 B/ray + 1 mask")
        Ri = 65536
        dpi, dgi, mi = dp_[:Ri].contiguous(), dg_[:Ri].contiguous(), mask[:Ri].contiguous()
        add("nnb_depth_invariant_fwd (2 radix selects + 2 passes, ONE block)", med_ms(lambda: _ops.DepthInvariant.apply(dpi, dgi, mi)), 16, Ri, "ray",
            "single-block kernel by design (a training batch is 4,096 rays): latency-bound, reported for completeness")
    # standalone compositing, forward and backward (3,088 / 4,624 B per ray)
    Rc = 262144            # 0.8 GB forward, 1.2 GB backward per launch: far beyond L2
    alpha = (torch.rand(Rc, 128, generator=g) * 0.2).to(dev).requires_grad_(True)
    rgb_s = torch.rand(Rc, 128, 3, generator=g).to(dev).requires_grad_(True)
    z = torch.sort(torch.rand(Rc, 128, generator=g) * 10)[0].to(dev)
    with torch.no_grad():
        add("nnb_composite_fwd (stage 3 standalone)", med_ms(lambda: _ops.Composite.apply(alpha, rgb_s, z, False, False)), 3088, Rc, "ray", "SURVEY 8d: 3,088 B/ray")
    rgb, dist, wts = _ops.Composite.apply(alpha, rgb_s, z, False, False)
    g_rgb, g_dist = torch.rand_like(rgb), torch.rand_like(dist)

    da, dc = torch.empty_like(alpha), torch.empty_like(rgb_s)
    from model._cabi import call, ptr, stream
    a_, c_ = alpha.detach(), rgb_s.detach()

    def bwd():     # the C-ABI call of Composite.backward without the autograd engine (it allocates per call)
        call("nnb_composite_bwd", ptr(a_), ptr(c_), ptr(z), ptr(wts.detach()), Rc, 128, 0, 0, ptr(g_rgb), ptr(g_dist), None, ptr(da), ptr(dc), stream())
    add("nnb_composite_bwd", med_ms(bwd), 4624, Rc, "ray", "SURVEY 8d: 4,624 B/ray")
    # reference-frame branch: back-projection and reprojection at 2 M points (h4 x w4 = 1024 x 2048)
    h4, w4 = 1024, 2048
    dfull = (torch.rand(1, 1, 4 * h4, 4 * w4, generator=g) + 1.0).to(dev)
    Km = K
    with torch.no_grad():
        pc = _ops.Backproject.apply(dfull, ss, Km, h4, w4, 0.01)
        add("nnb_backproject_fwd", med_ms(lambda: _ops.Backproject.apply(dfull, ss, Km, h4, w4, 0.01)), 4 + 12 + 1, h4 * w4, "point",
            "read one depth 4, write point 12 + clamp flag 1")
        i1 = torch.rand(3, h4, w4, generator=g).to(dev)
        i2 = torch.rand(3, h4, w4, generator=g).to(dev)
        Rm, tv = torch.eye(3, device=dev), torch.tensor([0.01, 0.0, -0.05], device=dev)
        add("nnb_reproj_fwd (stage 4 reprojection)", med_ms(lambda: _ops.Reproject.apply(i1, i2, pc, Rm, tv, Km, 0.01)), 64, h4 * w4, "point",
            "SURVEY 8d: 64 B/point")
        full = torch.rand(3, 4 * h4, 4 * w4, generator=g).to(dev)
        add("nnb_resize_bilinear (image -> point-cloud resolution)", med_ms(lambda: _ops.resize_bilinear(full, h4, w4)), 4 * 12 + 12, h4 * w4, "point",
            "4 taps x 12 B read + 12 B written per output pixel")
    return out


def fp16_mma_peak_tflops(dev):
    """Dense fp16 tensor throughput on this box (library GEMM, 8192^3, best of 5): the denominator of the Chamfer tensor fraction."""
    a = torch.randn(8192, 8192, device=dev, dtype=torch.float16)
    b = torch.randn(8192, 8192, device=dev, dtype=torch.float16)
    best = 1e9
    for _ in range(2):
        torch.matmul(a, b)
    for _ in range(5):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    del a, b
    return 2 * 8192 ** 3 / (best * 1e-3) / 1e12


def chamfer_rooflines(dev, sm_mhz, n_sm=148):
    """Stage 5 against its two bounds: the sweeps read every stage's 128 x 256 fp32 accumulator back through
    tcgen05.ld (64 pairs / clk / SM, two sweeps per direction: screen + exact), and SURVEY.md 8d's CUDA-core epilogue bound
    (3 fp32 lane-ops per pair over 128 lanes per SM).  Pairs are ALGORITHMIC pairs n_src x n_dst; stage lists may skip most of
    them on raster-ordered clouds, which is why that figure can exceed the sweep bounds."""
    from model import _ops
    clk = (sm_mhz or 1900.0) * 1e6
    ld_bound = 64.0 * n_sm * clk / 2.0           # pairs/s, two sweeps
    core_bound = 128.0 * n_sm * clk / 3.0
    fp16_peak = fp16_mma_peak_tflops(dev)
    g = torch.Generator().manual_seed(7)
    out = []

    def clouds(kind, n):
        if kind == "random":
            return (torch.rand(n, 3, generator=g) * 2 - 1).to(dev), (torch.rand(n, 3, generator=g) * 2 - 1).to(dev)
        hh = int(round((n * 250 / 400) ** 0.5))
        ww = n // hh
        v, u = torch.meshgrid(torch.linspace(-0.75, 0.75, hh), torch.linspace(-1, 1, ww), indexing="ij")
        zz = 2.0 + 0.3 * torch.sin(3 * u) * torch.cos(2 * v)
        a = torch.stack([u * zz, v * zz, -zz], -1).reshape(-1, 3)
        b = a + torch.tensor([0.013, -0.007, 0.02]) + 1e-3 * torch.randn(a.shape, generator=g)
        return a.to(dev).contiguous(), b.to(dev).contiguous()

    for kind, n in (("random", 28830), ("raster", 28830), ("random", 100000), ("raster", 100000)):
        a, b = clouds(kind, n)
        ms = med_ms(lambda: _ops.chamfer_nn(a, b), reps=7)
        pairs = float(a.shape[0]) * float(b.shape[0])
        pps = pairs / (ms * 1e-3)
        out.append({"kernel": "nnb_chamfer_nn (%s order, %d x %d)" % (kind, a.shape[0], b.shape[0]), "bound": "tcgen05.ld (TMEM read-back)",
                    "ms": ms, "unit": "T pairs/s", "achieved": pps / 1e12, "peak": ld_bound / 1e12, "frac": pps / ld_bound,
                    "cuda_core_epilogue_bound_tpairs": core_bound / 1e12, "frac_of_cuda_core_bound": pps / core_bound,
                    "algorithmic_gflop": 8 * pairs / 1e9,
                    "tensor_issued_tflops": 2 * 2 * 16 * pairs / (ms * 1e-3) / 1e12, "fp16_mma_peak_tflops_measured": fp16_peak,
                    "tensor_frac": (2 * 2 * 16 * pairs / (ms * 1e-3) / 1e12 / fp16_peak) if kind == "random" else None,
                    "note": "issued tensor work = 2 sweeps x one 128x256x16 fp16 MMA per 128 x 256 pairs (64 FLOP/pair) when no stage is "
                            "skipped (random order); raster order: the stage lists skip most pairs, so algorithmic pairs/s exceeds the "
                            "sweep bounds and the issued-work figure does not apply"})
        if kind == "raster":
            out[-1]["tensor_issued_tflops"] = None
    return out
