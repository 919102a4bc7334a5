"""Data-parallel plumbing for the hot path: one process per GPU,
samples sharded across ranks, no collective on the data path — coordinate maps,
kernel maps and features never leave their rank; only the (tiny) kernel-weight
gradients are summed.  Mirrors what the reference gets from
`torch.nn.parallel.DistributedDataParallel` around its example nets
(`examples/multigpu-ddp.py`, `ae_res.py` launcher) without wrapping the module.

Host-side only; works on any `torch.distributed` backend (NCCL on the GPUs,
gloo in the CPU tests)."""
from __future__ import annotations

from typing import Iterable, List, Sequence

import torch
import torch.distributed as dist


def shard_samples(sizes: Sequence[int], world: int) -> List[List[int]]:
    """Assign samples (by voxel count) to ranks: longest-processing-time greedy,
    deterministic (ties -> lower sample index, lower rank).  Every sample goes to
    exactly one rank; the heaviest rank is within one sample of the mean."""
    if world < 1:
        raise ValueError("world size must be >= 1")
    order = sorted(range(len(sizes)), key=lambda i: (-int(sizes[i]), i))
    load = [0] * world
    out: List[List[int]] = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda j: (load[j], j))
        out[r].append(i)
        load[r] += int(sizes[i])
    return [sorted(s) for s in out]


def allreduce_gradients(params: Iterable[torch.Tensor], group=None, average: bool = False) -> None:
    """Sum (or average) `.grad` of the given parameters over the ranks, in ONE
    flat bucket per dtype: the weights of a sparse conv net are a few MB, so the
    cost is launch latency, not bandwidth (NVSwitch: every GPU has full bandwidth
    to every peer), and one collective per dtype is the right size."""
    if not dist.is_available() or not dist.is_initialized() or dist.get_world_size(group) == 1:
        return
    by_dtype = {}
    for p in params:
        if p.grad is not None:
            by_dtype.setdefault(p.grad.dtype, []).append(p.grad)
    world = dist.get_world_size(group)
    for grads in by_dtype.values():
        flat = torch.cat([g.reshape(-1) for g in grads])
        dist.all_reduce(flat, group=group)
        if average:
            flat /= world
        off = 0
        for g in grads:
            n = g.numel()
            g.copy_(flat[off:off + n].view_as(g))
            off += n
