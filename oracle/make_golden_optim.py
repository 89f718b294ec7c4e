"""TEST INFRASTRUCTURE ONLY -- golden vector of the reference's learning-rate schedule.

Run in the build container (needs /root/reference):  python -m oracle.make_golden_optim
Imports the UNMODIFIED chemprop.nn_utils.NoamLR (nn_utils.py:115-194) through oracle/ref_shim.py, drives it the
way chemprop/train/train.py:82-86 does (optimizer.step(); scheduler.step()) and stores the learning rate seen after
construction and after every step in tests/golden/optim/noam_lr.npz.
"""
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle import ref_shim  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "optim", "noam_lr.npz")

CASES = {   # the reference's defaults (args.py: warmup_epochs 2.0, epochs 30, init/max/final lr 1e-4/1e-3/1e-4) and two others
    "default": dict(warmup_epochs=2.0, total_epochs=30, steps_per_epoch=25, init_lr=1e-4, max_lr=1e-3, final_lr=1e-4),
    "short": dict(warmup_epochs=0.5, total_epochs=3, steps_per_epoch=7, init_lr=5e-5, max_lr=2e-3, final_lr=1e-5),
    "past_the_end": dict(warmup_epochs=1, total_epochs=2, steps_per_epoch=4, init_lr=1e-4, max_lr=1e-3, final_lr=1e-4),
}


def main():
    ref_shim.load_reference()
    NoamLR = importlib.import_module("chemprop.nn_utils").NoamLR
    out = {}
    for name, c in CASES.items():
        w = torch.nn.Parameter(torch.zeros(3))
        opt = torch.optim.Adam([{"params": [w], "lr": c["init_lr"], "weight_decay": 0.0}])
        sch = NoamLR(optimizer=opt, warmup_epochs=[c["warmup_epochs"]], total_epochs=[c["total_epochs"]],
                     steps_per_epoch=c["steps_per_epoch"], init_lr=[c["init_lr"]], max_lr=[c["max_lr"]],
                     final_lr=[c["final_lr"]])
        n = c["total_epochs"] * c["steps_per_epoch"] + 3
        lrs, steps = [opt.param_groups[0]["lr"]], [sch.current_step]
        for _ in range(n):
            w.grad = torch.ones(3)
            opt.step()
            sch.step()
            lrs.append(opt.param_groups[0]["lr"])
            steps.append(sch.current_step)
        out[name + "/lr"] = np.array(lrs, dtype=np.float64)
        out[name + "/current_step"] = np.array(steps, dtype=np.int64)
        out[name + "/cfg"] = np.array([c["warmup_epochs"], c["total_epochs"], c["steps_per_epoch"], c["init_lr"], c["max_lr"],
                                       c["final_lr"]], dtype=np.float64)
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    np.savez(OUT, **out)
    print("wrote", OUT, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
