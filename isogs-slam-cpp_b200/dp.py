"""Keyframe-sharded data-parallel mapping: host-side plumbing only.

Every rank holds the full (replicated) Gaussian map and Adam state, renders a DIFFERENT keyframe of
the selected window per iteration, the per-parameter gradients live in ONE flat fp32 buffer, a single
all-reduce (NCCL over NVLink/NVSwitch on the GPU box, gloo in the CPU tests) averages it, and every
rank applies the identical Adam step -- replicas stay bit-identical without any broadcast.  The
reference has no distributed code; its sequential analogue is one uniformly drawn keyframe
per iteration (python_ordinary/scripts/splatam.py:1415-1426).
"""
import torch
import torch.distributed as dist


def keyframe_for(iteration, rank, world, window):
    """Keyframe slot rendered by `rank` at `iteration`: `world` different draws from the window per
    iteration, round-robin so that every rank sees every keyframe."""
    return (iteration * world + rank) % window


def allreduce_mean_(flat, group=None):
    """In-place average of the flat gradient buffer over all ranks (sum all-reduce, then 1/world).
    Returns the work-free tensor; a single collective covers every parameter group."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return flat
    if dist.get_backend(group) == "nccl":
        dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=group)   # ncclAvg: the 1/world scaling rides in the collective
    else:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        flat.mul_(1.0 / world)
    return flat


def replicas_identical(tensors, group=None):
    """True on every rank iff the given tensors are bit-identical across ranks (a debugging / test aid:
    compares a 64-bit checksum of the raw bits with max and min all-reduces)."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return True
    sums = []
    for t in tensors:
        b = t.detach().contiguous().view(torch.int32).to(torch.int64)
        w = torch.arange(1, b.numel() + 1, dtype=torch.int64, device=b.device)
        sums.append(((b.reshape(-1) * w) % 2147483629).sum())
    s = torch.stack(sums)
    hi, lo = s.clone(), s.clone()
    dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
    dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
    return bool(torch.equal(hi, lo))
