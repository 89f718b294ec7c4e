"""Fused optimiser step for the model's parameters.

Drop-ins for the two objects the reference's training loop steps after ``loss.backward()``
(chemprop/train/train.py:79-86):

* ``FlatAdam``  -- ``torch.optim.Adam`` / ``AdamW`` as ``build_optimizer`` creates them
  (chemprop/utils.py:295-310: one parameter group, ``lr = init_lr``, ``weight_decay``), plus the optional
  ``clip_grad_norm_`` of train.py:80-81.  The parameters are re-pointed at views of ONE flat fp32 buffer, the
  moments are flat buffers too, and ``step()`` is one call of ``wd_adam_step`` (include/wdmpnn.h): one HBM pass
  instead of torch's ~10 for-each kernels, and the gradient buffer is the same flat buffer the data-parallel
  all-reduce fills (``dp.FlatGradAllReduce``), so nothing is copied between backward, all-reduce and update.
* ``NoamLR``    -- chemprop/nn_utils.py:115-194 (linear warm-up, exponential decay), same constructor, same
  ``step()`` / ``get_lr()``; it writes ``param_groups[i]['lr']`` exactly as the reference's does, so either
  scheduler drives either optimiser.

There is no CPU path: ``FlatAdam`` raises for parameters that are not on a CUDA device."""
import ctypes as C
from typing import Iterable, List, Optional, Union

import numpy as np
import torch

from . import _capi


class FlatAdam(torch.optim.Optimizer):
    """Adam / AdamW over one flat parameter buffer.  ``decoupled=True`` is AdamW; ``max_grad_norm`` fuses
    ``clip_grad_norm_(params, max_grad_norm)`` into the step; ``grad_buffer``: a ``dp.FlatGradAllReduce`` over
    the same parameters in the same order, whose flat (already reduced) buffer is then read directly."""

    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, decoupled: bool = False, max_grad_norm: Optional[float] = None,
                 grad_buffer=None):
        params = list(params)
        if params and isinstance(params[0], dict):
            if len(params) != 1:
                raise ValueError("FlatAdam takes one parameter group (chemprop/utils.py:304 builds exactly one)")
            group = dict(params[0])
            params = list(group.pop("params"))
            lr = group.pop("lr", lr)
            weight_decay = group.pop("weight_decay", weight_decay)
            betas, eps = group.pop("betas", betas), group.pop("eps", eps)
        params = [p for p in params if p.requires_grad]
        if not params:
            raise ValueError("no trainable parameters")
        dev = params[0].device
        if dev.type != "cuda":
            raise RuntimeError("FlatAdam runs wd_adam_step on a CUDA device; there is no CPU fallback")
        if any(p.device != dev or p.dtype != torch.float32 for p in params):
            raise ValueError("all parameters must be fp32 tensors on one CUDA device")
        if not 0.0 <= betas[0] < 1.0 or not 0.0 <= betas[1] < 1.0 or lr < 0 or eps < 0 or weight_decay < 0:
            raise ValueError("invalid Adam hyper-parameters")
        super().__init__(params, dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay))
        self.decoupled, self.max_grad_norm = bool(decoupled), max_grad_norm
        self._params = params
        self._sizes = [p.numel() for p in params]
        n = sum(self._sizes)
        self.flat_params = torch.empty(n, device=dev, dtype=torch.float32)
        for p, v in zip(params, self.flat_params.split(self._sizes)):
            v.copy_(p.data.reshape(-1))
            p.data = v.view(p.shape)                       # the module now reads the flat buffer
        self.exp_avg = torch.zeros(n, device=dev, dtype=torch.float32)
        self.exp_avg_sq = torch.zeros(n, device=dev, dtype=torch.float32)
        self._grad_buffer = grad_buffer
        if grad_buffer is not None:
            if [id(p) for p in grad_buffer.params] != [id(p) for p in params]:
                raise ValueError("grad_buffer must cover the same parameters in the same order")
            self.flat_grads = grad_buffer.flat
        else:
            self.flat_grads = torch.zeros(n, device=dev, dtype=torch.float32)
        self._grad_views = list(self.flat_grads.split(self._sizes))
        self._scratch = torch.zeros(int(_capi.lib().wd_adam_scratch_floats()), device=dev, dtype=torch.float32)
        self.grad_norm = torch.zeros(1, device=dev, dtype=torch.float32)   # ||g||_2 of the last clipped step
        self.steps = 0
        for p, m, v in zip(params, self.exp_avg.split(self._sizes), self.exp_avg_sq.split(self._sizes)):
            self.state[p] = {"step": torch.tensor(0.0), "exp_avg": m.view(p.shape), "exp_avg_sq": v.view(p.shape)}

    def _pack_grads(self):
        """Returns the parameters WITHOUT a gradient this step.  torch's Adam skips those entirely (no weight decay, no
        moment decay, no update: e.g. W_h of a depth-1 encoder, which the reference's graph never touches), so their
        slices of the parameter and moment buffers are saved here and put back after the fused step."""
        skipped = [i for i, p in enumerate(self._params) if p.grad is None]
        if self._grad_buffer is not None:
            return skipped                                  # FlatGradAllReduce.__call__ has filled (and reduced) it
        src, dst = [], []
        for p, v in zip(self._params, self._grad_views):
            if p.grad is None:
                v.zero_()
            elif p.grad.data_ptr() != v.data_ptr():
                src.append(p.grad.reshape(-1))
                dst.append(v)
                p.grad = v.view(p.shape)                    # zero_grad(set_to_none=False) keeps the flat views alive
        if src:
            torch._foreach_copy_(dst, src)                  # one fused launch for all parameters
        return skipped

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        skipped = self._pack_grads()
        keep = []
        if skipped:                                         # rare (frozen-by-absence parameters): save, step, restore
            offs = np.concatenate([[0], np.cumsum(self._sizes)])
            for i in skipped:
                sl = slice(int(offs[i]), int(offs[i + 1]))
                keep.append((sl, self.flat_params[sl].clone(), self.exp_avg[sl].clone(), self.exp_avg_sq[sl].clone()))
        g = self.param_groups[0]
        self.steps += 1
        cfg = _capi.WdAdam(float(g["lr"]), float(g["betas"][0]), float(g["betas"][1]), float(g["eps"]),
                           float(g["weight_decay"]), float(self.max_grad_norm) if self.max_grad_norm else 0.0,
                           self.steps, int(self.decoupled))
        p = _capi.ptr
        with _capi.device_guard(self.flat_params.device):    # the library launches on the current device
            _capi.check(_capi.lib().wd_adam_step(p(self.flat_params), p(self.flat_grads), p(self.exp_avg), p(self.exp_avg_sq),
                                                 self.flat_params.numel(), C.byref(cfg), p(self._scratch), p(self.grad_norm),
                                                 _capi.current_stream(self.flat_params.device)))
        for sl, pv, m, v in keep:
            self.flat_params[sl].copy_(pv)
            self.exp_avg[sl].copy_(m)
            self.exp_avg_sq[sl].copy_(v)
        return loss

    def state_dict(self):
        for st in self.state.values():                      # torch.optim.Adam's layout: a step count per parameter
            st["step"] = torch.tensor(float(self.steps))
        return super().state_dict()

    def load_state_dict(self, state_dict):
        super().load_state_dict(state_dict)
        steps = 0
        for p, m, v in zip(self._params, self.exp_avg.split(self._sizes), self.exp_avg_sq.split(self._sizes)):
            st = self.state.get(p, {})
            if "exp_avg" in st:
                m.copy_(st["exp_avg"].reshape(-1))
                v.copy_(st["exp_avg_sq"].reshape(-1))
                steps = max(steps, int(float(st.get("step", 0))))
            self.state[p] = {"step": torch.tensor(float(steps)), "exp_avg": m.view(p.shape), "exp_avg_sq": v.view(p.shape)}
        self.steps = steps


def noam_lr_at(step: int, warmup_steps: int, total_steps: int, init_lr: float, max_lr: float, final_lr: float) -> float:
    """Closed form of the schedule chemprop/nn_utils.py:115-194 walks step by step: linear from ``init_lr`` to ``max_lr``
    over ``warmup_steps``, then geometric down to ``final_lr`` at ``total_steps``, constant afterwards.  float64
    scalars in the same operation order as the reference, so the values agree to the last bit (golden-pinned)."""
    init_lr, max_lr, final_lr = np.float64(init_lr), np.float64(max_lr), np.float64(final_lr)
    if step <= warmup_steps:
        return float(init_lr + step * ((max_lr - init_lr) / warmup_steps))
    if step <= total_steps:
        return float(max_lr * (((final_lr / max_lr) ** (1 / (total_steps - warmup_steps))) ** (step - warmup_steps)))
    return float(final_lr)


class NoamLR:
    """Per-step scheduler with the reference NoamLR's constructor and ``step()`` / ``get_lr()`` surface, as a thin
    driver of ``noam_lr_at``.  (FlatAdam is a torch.optim.Optimizer, so the reference's own NoamLR drives it unchanged;
    this class exists for callers that run without an importable chemprop.)"""

    def __init__(self, optimizer, warmup_epochs: List[Union[float, int]], total_epochs: List[int], steps_per_epoch: int,
                 init_lr: List[float], max_lr: List[float], final_lr: List[float]):
        groups = optimizer.param_groups
        if not all(len(x) == len(groups) for x in (warmup_epochs, total_epochs, init_lr, max_lr, final_lr)):
            raise ValueError("one entry per parameter group is required")
        self.optimizer, self.steps_per_epoch = optimizer, steps_per_epoch
        self._plan = [(int(np.float64(w) * steps_per_epoch), t * steps_per_epoch, i, m, f)
                      for w, t, i, m, f in zip(warmup_epochs, total_epochs, init_lr, max_lr, final_lr)]
        self.current_step = 0
        self.lr = [float(i) for i in init_lr]
        for g in groups:
            g.setdefault("initial_lr", g["lr"])
        self.step()      # torch's _LRScheduler.__init__ (the reference's base class) steps once: current_step == 1 now

    def get_lr(self) -> List[float]:
        return list(self.lr)

    def step(self, current_step: int = None):
        self.current_step = self.current_step + 1 if current_step is None else current_step
        for i, (group, plan) in enumerate(zip(self.optimizer.param_groups, self._plan)):
            self.lr[i] = group["lr"] = noam_lr_at(self.current_step, *plan)
