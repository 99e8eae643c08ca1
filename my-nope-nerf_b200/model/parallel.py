"""Data parallelism for the hot path: one process per GPU, ray batches (training) or pixel bands (rendering) sharded
by rank, and ONE flat-buffer gradient allreduce per step over NCCL/NVLink (gloo in CPU tests).  The reference has no
multi-GPU code; the 1-GPU result on the same global batch is the parity oracle."""
import os

import torch
import torch.distributed as dist


class DataParallelContext(object):
    def __init__(self, rank=0, world_size=1, group=None):
        self.rank, self.world_size, self.group = rank, world_size, group
        self.timing = None
        self.bytes_per_allreduce = 0
        self.check_symmetry = os.environ.get('NNB_DP_CHECK', '0') == '1'

    @classmethod
    def from_env(cls):
        if dist.is_available() and dist.is_initialized():
            return cls(dist.get_rank(), dist.get_world_size())
        return cls(0, 1)

    def shard(self, n):
        """Contiguous [lo, hi) slice of n units for this rank (first n % world ranks take one extra)."""
        return shard_range(n, self.rank, self.world_size)

    def allreduce_grads(self, params, extra=()):
        """Average the gradients of `params` across ranks with ONE collective over a flat fp32 buffer.  Parameters whose grad
        is None occupy a zero slot (so every rank's buffer has the same layout) and stay without gradient afterwards.  `extra`: scalar tensors (the loss terms of
        the step) averaged in place by the same collective.  `self.timing` = [] collects CUDA-event pairs around the call."""
        if self.world_size == 1:
            return
        params = [p for p in params if p.requires_grad]
        if not params:
            return
        ev = None
        if self.timing is not None and torch.cuda.is_available():
            ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            ev[0].record()
        extra = [e for e in extra if torch.is_tensor(e)]
        if self.check_symmetry:
            # debugging aid (env NNB_DP_CHECK=1; one extra collective and a host read): every rank must reach the same parameters
            have = torch.tensor([float(p.grad is not None) for p in params], device=params[0].device)
            lo, hi = have.clone(), have.clone()
            dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=self.group)
            dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=self.group)
            if not torch.equal(lo, hi):
                raise RuntimeError("data-parallel ranks disagree on which parameters received a gradient: %s"
                                   % [i for i, (a, b) in enumerate(zip(lo.tolist(), hi.tolist())) if a != b])
        flat = flatten_grads(params, extra)
        if dist.get_backend(self.group) == 'nccl':
            dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=self.group)     # the average is taken inside the collective
        else:
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
            flat.div_(self.world_size)
        unflatten_grads(flat, params, extra)
        if ev is not None:
            ev[1].record()
            self.timing.append(ev)
            self.bytes_per_allreduce = flat.numel() * 4

    def gather_rows(self, local, n_total):
        """All ranks' row shards (in rank order) concatenated on every rank: render bands -> full frame."""
        if self.world_size == 1:
            return local
        sizes = [shard_range(n_total, r, self.world_size) for r in range(self.world_size)]
        maxn = max(hi - lo for lo, hi in sizes)
        pad = torch.zeros((maxn,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
        pad[:local.shape[0]] = local
        outs = [torch.empty_like(pad) for _ in range(self.world_size)]
        dist.all_gather(outs, pad, group=self.group)
        return torch.cat([o[:hi - lo] for o, (lo, hi) in zip(outs, sizes)], dim=0)


def shard_range(n, rank, world_size):
    base, rem = divmod(n, world_size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def flatten_grads(params, extra=()):
    """One fp32 buffer holding every gradient (zeros where a parameter has no gradient on this rank) followed by the `extra`
    scalars: one `cat` launch."""
    return torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1).float() for p in params]
                     + [e.detach().reshape(-1).float() for e in extra])


def unflatten_grads(flat, params, extra=()):
    """Inverse of flatten_grads.  A parameter without a gradient keeps none: which parameters a step reaches is a property of
    the step (frame pairing, active loss terms), the same on every rank, and the optimisers treat "no gradient" (skipped)
    differently from a zero one (moments still move) -- as torch.optim.Adam does under the reference."""
    views, dst, o = [], [], 0
    for p in params:
        n = p.numel()
        if p.grad is not None:
            views.append(flat[o:o + n].view_as(p))
            dst.append(p.grad)
        o += n
    for e in extra:
        n = e.numel()
        views.append(flat[o:o + n].view_as(e))
        dst.append(e.detach())
        o += n
    with torch.no_grad():
        torch._foreach_copy_(dst, views)


class AllGatherRows(torch.autograd.Function):
    """Every rank's shard of a 1-D tensor concatenated in rank order (shards as DataParallelContext.shard cuts them).  The
    caller evaluates the same function of the gathered tensor on every rank and scales it by the world size, so the backward
    is the slice of the incoming gradient that belongs to this rank's shard: no collective in the backward pass."""

    @staticmethod
    def forward(ctx, local, dp, n_total):
        ctx.range = dp.shard(n_total)
        return dp.gather_rows(local.detach().contiguous(), n_total)

    @staticmethod
    def backward(ctx, g):
        lo, hi = ctx.range
        return g[lo:hi].contiguous(), None, None


def init_from_env(backend=None):
    """Initialise torch.distributed from torchrun's environment (RANK/LOCAL_RANK/WORLD_SIZE/MASTER_*)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world <= 1:
        return DataParallelContext(0, 1)
    if not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group(backend=backend)
    return DataParallelContext(dist.get_rank(), dist.get_world_size())
