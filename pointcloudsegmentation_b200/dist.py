"""Block sharding + the one collective on the path.

The reference runs one tower per GPU inside one process, gives each tower its own block
(train_graph_pool.py:95-108), pads the batch to a multiple of num_gpus by repeating items
(provider.py:25-40) and averages tower gradients per variable on the CPU (train_util.py:77-111).
Here: one process per GPU (torch.distributed, NCCL over NVLink/NVSwitch; gloo in CPU tests), blocks
dealt to ranks by edge-count-balanced greedy assignment, and ONE all-reduce of the flattened fp32
weight gradient per step.  Every custom op is local to a block, so nothing else crosses GPUs;
inference shards blocks with no communication at all.
"""
from __future__ import annotations

from typing import Iterable, Sequence

import torch
import torch.distributed as dist


def shard_blocks(costs: Sequence[float], world: int) -> list[list[int]]:
    """Deal block indices to `world` ranks, heaviest first onto the lightest rank (edge count, not
    point count, predicts time).  Every rank gets at least one block when len(costs) >= world;
    when there are fewer blocks than ranks the list is padded by repeating blocks, like
    provider.py:25-40 pads a batch to a multiple of num_gpus."""
    n = len(costs)
    if n == 0:
        return [[] for _ in range(world)]
    order = sorted(range(n), key=lambda i: (-float(costs[i]), i))
    if n < world:
        order = order + [order[i % n] for i in range(world - n)]
    loads = [0.0] * world
    out: list[list[int]] = [[] for _ in range(world)]
    for i in order:
        r = min(range(world), key=lambda k: (loads[k], len(out[k]), k))
        out[r].append(i)
        loads[r] += float(costs[i])
    return [sorted(x) for x in out]


def fill_to_budget(sizes: Sequence[int], n: int, target: int) -> list[int]:
    """Pick `n` of the candidate blocks (indices into `sizes`, ascending) whose sizes add up as close to `target` as
    single swaps get: every rank of a synchronous data-parallel step should carry the same work, because the step takes
    as long as its heaviest rank (the straggler cost grows with the number of ranks; SCALE_r01: 3.5 % at 8 GPUs with
    unbalanced random batches).  Start from the first n candidates, then repeatedly apply the swap (one in, one out)
    that brings the total closest to the target."""
    m = len(sizes)
    assert 0 < n <= m
    chosen = list(range(n))
    rest = list(range(n, m))
    total = sum(int(sizes[i]) for i in chosen)
    while rest:
        best = None
        for a, i in enumerate(chosen):
            for b, j in enumerate(rest):
                d = abs(total - int(sizes[i]) + int(sizes[j]) - target)
                if d < abs(total - target) and (best is None or d < best[0]):
                    best = (d, a, b)
        if best is None:
            break
        _, a, b = best
        total += int(sizes[rest[b]]) - int(sizes[chosen[a]])
        chosen[a], rest[b] = rest[b], chosen[a]
    return sorted(chosen)


class GradAverager:
    """Mean of weight gradients over ranks through a single flattened bucket (these networks have
    ~1-5 M parameters: one 7-20 MB all-reduce, latency-bound on NVLink 5, so one call not many)."""

    def __init__(self, params: Iterable[torch.nn.Parameter], group=None):
        self.params = [p for p in params]
        self.group = group
        n = sum(p.numel() for p in self.params)
        p0 = self.params[0]
        self.flat = torch.zeros(n, dtype=p0.dtype, device=p0.device)
        self.views = [v.view_as(p) for v, p in zip(self.flat.split([p.numel() for p in self.params]), self.params)]

    def __call__(self) -> None:
        if not dist.is_initialized() or dist.get_world_size(self.group) == 1:
            return
        world = dist.get_world_size(self.group)
        have = [i for i, p in enumerate(self.params) if p.grad is not None]
        if len(have) != len(self.params):  # a parameter unused this step contributes zeros
            self.flat.zero_()
        # multi-tensor copies: two fused launches instead of two per parameter
        torch._foreach_copy_([self.views[i] for i in have], [self.params[i].grad for i in have])
        if dist.get_backend(self.group) == "nccl":
            dist.all_reduce(self.flat, op=dist.ReduceOp.AVG, group=self.group)
        else:
            dist.all_reduce(self.flat, group=self.group)
            self.flat.div_(world)
        for i, p in enumerate(self.params):
            if p.grad is None:
                p.grad = torch.empty_like(p)
        torch._foreach_copy_([p.grad for p in self.params], self.views)


def broadcast_parameters(params: Iterable[torch.nn.Parameter], src: int = 0, group=None) -> None:
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        for p in params:
            dist.broadcast(p.data, src, group=group)
