"""Data-parallel plumbing for the encoder.

The reference has no multi-GPU support at all.  Graphs in a BatchMolGraph are a disjoint union
(chemprop/features/featurization.py:782-800: per-molecule index offsets, no cross-molecule edge),
so the path shards by molecule with NO data-path collective; the only exchange is one all-reduce of
the weight gradients per step.  One process per GPU (torchrun), NCCL over NVLink / NVSwitch; the same
code runs on gloo for the CPU tests.

The gradient payload is tiny at h=300 (264 300 floats = 1.06 MB) and latency-bound, so all parameter
gradients are packed into ONE contiguous buffer and reduced with a single collective.
"""
from typing import Iterable, List, Optional, Sequence

import torch
import torch.distributed as dist


def shard_range(n_items: int, rank: int, world: int) -> range:
    """Contiguous, equal-sized shard of ``n_items`` for ``rank`` (remainder dropped so that every rank
    holds the same number of molecules: the reference normalises the loss by the local batch,
    chemprop/train/train.py:74, and mean-of-means equals the global mean only for equal shards)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"invalid rank {rank} / world {world}")
    per = n_items // world
    return range(rank * per, (rank + 1) * per)


def shard_graphs(graphs: Sequence, rank: int, world: int) -> List:
    return [graphs[i] for i in shard_range(len(graphs), rank, world)]


class FlatGradAllReduce:
    """Packs the ``.grad`` of ``params`` into one flat fp32 buffer, all-reduces it once and scatters the
    (averaged) result back into the ``.grad`` tensors.  Parameters without a gradient contribute zeros
    (every rank must issue the same collective)."""

    def __init__(self, params: Iterable[torch.nn.Parameter], group: Optional[dist.ProcessGroup] = None,
                 average: bool = True):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("no trainable parameters")
        self.group = group
        self.average = average
        self.sizes = [p.numel() for p in self.params]
        dev, dt = self.params[0].device, self.params[0].dtype
        if any(p.device != dev or p.dtype != dt for p in self.params):
            raise ValueError("all parameters must share one device and dtype")
        self.flat = torch.zeros(sum(self.sizes), device=dev, dtype=dt)
        self.views = list(self.flat.split(self.sizes))

    @property
    def nbytes(self) -> int:
        return self.flat.numel() * self.flat.element_size()

    def __call__(self) -> torch.Tensor:
        world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        for p, v in zip(self.params, self.views):
            if p.grad is None:
                v.zero_()
            else:
                v.copy_(p.grad.reshape(-1))
        if world > 1:
            dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
            if self.average:
                self.flat.div_(world)
        for p, v in zip(self.params, self.views):
            if p.grad is None:
                p.grad = v.view_as(p).clone()
            else:
                p.grad.copy_(v.view_as(p))
        return self.flat
