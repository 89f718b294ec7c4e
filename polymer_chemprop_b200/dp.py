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
    """One flat fp32 buffer for the gradients of ``params``, all-reduced with ONE collective per step.

    ``attach(model)``: every ``MPNEncoder`` of ``model`` whose parameters are all in ``params`` writes its weight
    gradients STRAIGHT into this buffer (``wd_encoder_backward`` takes plain pointers: ``WdGrads`` are views of
    ``flat``) and each such parameter's ``.grad`` IS its view, so nothing is packed or unpacked around the
    collective: per step the data-parallel path costs exactly one ``ncclAllReduce(avg)``.  Parameters outside an
    attached encoder (an FFN head, say) are packed / unpacked with one fused copy each way.  Parameters without a
    gradient contribute zeros (every rank must issue the same collective)."""

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
        self._view_of = {id(p): v.view_as(p) for p, v in zip(self.params, self.views)}
        self._attached = {}         # id(encoder) -> (encoder, [views in WdGrads order], [params])
        self._written = set()       # encoders whose backward has written this step's gradients into the buffer

    @property
    def nbytes(self) -> int:
        return self.flat.numel() * self.flat.element_size()

    # ---- gradient sink of the encoders (polymer_chemprop_b200.mpn._EncoderFunction.backward) -----------------
    def attach(self, model: torch.nn.Module) -> int:
        """Makes this buffer the gradient sink of the MPNEncoders in ``model``; returns how many were attached."""
        from .mpn import MPNEncoder
        n = 0
        for enc in {id(m): m for m in model.modules() if isinstance(m, MPNEncoder)}.values():
            ps = (enc.W_i.weight, enc.W_i.bias, enc.W_h.weight, enc.W_h.bias, enc.W_o.weight, enc.W_o.bias)
            if any(p is not None and id(p) not in self._view_of for p in ps):
                continue
            self._attached[id(enc)] = (enc, [self._view_of[id(p)] if p is not None else None for p in ps], ps)
            enc._grad_sink = self
            n += 1
        return n

    def detach(self):
        for enc, _v, _p in self._attached.values():
            enc._grad_sink = None
        self._attached.clear()

    def take(self, enc) -> bool:
        """True: the calling backward may OVERWRITE the encoder's slices (its first backward since the last
        all-reduce); False: a second application of a shared encoder -- it must go through autograd's accumulation."""
        if id(enc) not in self._attached or id(enc) in self._written:
            return False
        _e, views, ps = self._attached[id(enc)]
        if any(p is not None and p.grad is not None and p.grad.data_ptr() != v.data_ptr() for p, v in zip(ps, views)):
            return False                                    # someone else has already accumulated into a separate .grad
        self._written.add(id(enc))
        for p, v in zip(ps, views):
            if p is not None:
                p.grad = v
        return True

    def views_for(self, enc):
        return self._attached[id(enc)][1]

    def __call__(self) -> torch.Tensor:
        world = dist.get_world_size(self.group) if dist.is_initialized() else 1
        src, dst = [], []
        for p, v in zip(self.params, self.views):
            if p.grad is None:
                v.zero_()
            elif p.grad.data_ptr() != v.data_ptr():
                src.append(p.grad.reshape(-1))
                dst.append(v)
        if src:
            torch._foreach_copy_(dst, src)                  # one fused launch for everything outside the sinks
        if world > 1:
            backend = dist.get_backend(self.group)
            if self.average and backend == "nccl":          # ncclAvg: the 1/R scale rides inside the collective
                dist.all_reduce(self.flat, op=dist.ReduceOp.AVG, group=self.group)
            else:
                dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
                if self.average:
                    self.flat.div_(world)
        for p in self.params:                               # every .grad is now its view of the reduced buffer
            v = self._view_of[id(p)]
            if p.grad is None or p.grad.data_ptr() != v.data_ptr():
                p.grad = v
        self._written.clear()
        return self.flat
