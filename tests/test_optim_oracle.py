"""CPU: the optimiser oracle (oracle/optim_oracle.py) against torch.optim.Adam / AdamW + clip_grad_norm_ run live
(the third-party code the reference calls, chemprop/utils.py:295-310, chemprop/train/train.py:80-82), the NoamLR
restatement and the product's NoamLR mirror against the golden vector recorded from the unmodified reference class
(tests/golden/optim/noam_lr.npz, oracle/make_golden_optim.py), and the host-side checks of FlatAdam."""
import os

import numpy as np
import pytest
import torch

from oracle import optim_oracle
from polymer_chemprop_b200 import FlatAdam, NoamLR

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "optim", "noam_lr.npz")
TOL = 2e-6      # fp32 updates, different association of the same formula


@pytest.mark.parametrize("kind,wd,clip", [("adam", 0.0, None), ("adam", 0.01, None), ("adamw", 0.05, None),
                                          ("adam", 0.0, 0.5), ("adamw", 0.01, 2.0)])
def test_adam_oracle_matches_torch(kind, wd, clip):
    rng = np.random.default_rng(3)
    shapes = [(7, 5), (5,), (3, 4)]
    params = [torch.nn.Parameter(torch.from_numpy(rng.standard_normal(s).astype(np.float32))) for s in shapes]
    cls = torch.optim.AdamW if kind == "adamw" else torch.optim.Adam
    opt = cls([{"params": params, "lr": 3e-3, "weight_decay": wd}])
    p = np.concatenate([q.detach().numpy().ravel() for q in params])
    m, v = np.zeros_like(p), np.zeros_like(p)
    for step in range(1, 13):
        grads = [rng.standard_normal(s).astype(np.float32) * (1.0 if step % 3 else 4.0) for s in shapes]
        for q, g in zip(params, grads):
            q.grad = torch.from_numpy(g.copy())
        if clip:
            torch.nn.utils.clip_grad_norm_(params, clip)
        lr = 3e-3 * (1 + 0.1 * step)                       # a scheduler moves the rate every step
        opt.param_groups[0]["lr"] = lr
        opt.step()
        p, m, v = optim_oracle.adam_step(p, np.concatenate([g.ravel() for g in grads]), m, v, step, lr=lr,
                                         weight_decay=wd, decoupled=kind == "adamw", max_grad_norm=clip)
        ref = np.concatenate([q.detach().numpy().ravel() for q in params])
        assert np.max(np.abs(p - ref)) <= TOL * max(1.0, np.max(np.abs(ref))), step


def _cases():
    d = np.load(GOLDEN)
    for name in sorted({k.split("/")[0] for k in d.files}):
        w, t, s, i, mx, f = d[name + "/cfg"]
        yield name, dict(warmup_epochs=float(w), total_epochs=int(t), steps_per_epoch=int(s), init_lr=float(i),
                         max_lr=float(mx), final_lr=float(f)), d[name + "/lr"], d[name + "/current_step"]


def test_noam_oracle_matches_reference_golden():
    n = 0
    for name, c, lrs, steps in _cases():
        for lr, s in zip(lrs, steps):
            assert optim_oracle.noam_lr(int(s), **c) == pytest.approx(lr, rel=1e-14, abs=0), (name, s)
            n += 1
    assert n > 700


class _Groups:                  # what NoamLR needs from an optimiser
    def __init__(self, lr):
        self.param_groups = [{"lr": lr}]


def test_noam_mirror_matches_reference_golden_step_by_step():
    """Same constructor, same calling sequence as chemprop/train/train.py:82-86; equal to the last bit."""
    for name, c, lrs, steps in _cases():
        opt = _Groups(c["init_lr"])
        sch = NoamLR(opt, [c["warmup_epochs"]], [c["total_epochs"]], c["steps_per_epoch"], [c["init_lr"]], [c["max_lr"]],
                     [c["final_lr"]])
        got, cur = [opt.param_groups[0]["lr"]], [sch.current_step]
        for _ in range(len(lrs) - 1):
            sch.step()
            got.append(opt.param_groups[0]["lr"])
            cur.append(sch.current_step)
        assert cur == list(steps), name
        assert np.array_equal(np.array(got), lrs), name
        assert sch.get_lr() == [got[-1]]


def test_noam_mirror_rejects_mismatched_lists():
    with pytest.raises(ValueError):
        NoamLR(_Groups(1e-4), [2.0, 1.0], [30], 10, [1e-4], [1e-3], [1e-4])


def test_flat_adam_has_no_cpu_path():
    w = torch.nn.Parameter(torch.zeros(4))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        FlatAdam([w])
    with pytest.raises(ValueError):
        FlatAdam([])
