"""Data parallelism for the hot path: one process per GPU, ray batches (training) or pixel bands (rendering) sharded
by rank, and ONE flat-buffer gradient allreduce per step over NCCL/NVLink (gloo in CPU tests).  The reference has no
multi-GPU code; the 1-GPU result on the same global batch is the parity oracle."""
import os

import torch
import torch.distributed as dist


class DataParallelContext(object):
    def __init__(self, rank=0, world_size=1, group=None):
        self.rank, self.world_size, self.group = rank, world_size, group

    @classmethod
    def from_env(cls):
        if dist.is_available() and dist.is_initialized():
            return cls(dist.get_rank(), dist.get_world_size())
        return cls(0, 1)

    def shard(self, n):
        """Contiguous [lo, hi) slice of n units for this rank (first n % world ranks take one extra)."""
        return shard_range(n, self.rank, self.world_size)

    def allreduce_grads(self, params):
        """Average the gradients of `params` across ranks with one collective over a flat fp32 buffer.
        Parameters whose grad is None on this rank contribute zeros (every rank must pass the same list)."""
        if self.world_size == 1:
            return
        params = [p for p in params if p.requires_grad]
        if not params:
            return
        flat = flatten_grads(params)
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
        flat.div_(self.world_size)
        unflatten_grads(flat, params)

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


def flatten_grads(params):
    """One fp32 buffer holding every gradient (zeros where a parameter has no gradient on this rank): one `cat` launch."""
    return torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1).float() for p in params])


def unflatten_grads(flat, params):
    views, o = [], 0
    for p in params:
        n = p.numel()
        views.append(flat[o:o + n].view_as(p))
        o += n
    missing = [i for i, p in enumerate(params) if p.grad is None]
    for i in missing:
        params[i].grad = torch.empty_like(params[i])
    torch._foreach_copy_([p.grad for p in params], views)


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
