"""GPU: the drop-in encoder inside the reference's FULL model (encoder + FFN head + masked MSE loss).

tests/golden/model/*.npz hold the state_dict, predictions, loss and parameter gradients of the UNMODIFIED
chemprop.models.model.MoleculeModel run by oracle/make_golden_model.py (forward as model.py:174-194, loss as
train.py:46-74).  Here the same state_dict is loaded into this package's MPN plus the FFN head of
model.py:57-116 (plain torch layers: the head is outside the accelerated path and consumes the encoder output
unchanged).  Tolerances: BASELINE.json's north star asks <= 1e-4 relative on predictions for a tensor-core
path.  MEASURED on B200: the FP32 path is at 2e-7 .. 5e-7 (asserted <= 1e-5); the default tensor-core mode, tf32x3
(error-compensated forward contractions, tc_gemm4.cuh), measures 1.7e-6 and 1.1e-5 and is asserted at the north
star's 1e-4; the single-pass TF32 mode is at 1.5e-4 (h=64, 3 tasks) and 5.3e-4 (h=300, 1 task:
predictions of magnitude ~0.3 out of cancelling sums; it rounds both operands to 11 bits, 2^-11 = 4.9e-4 per
element, the same order the numpy TF32 emulation of the reference arithmetic gives,
tests/test_oracle_golden.py::test_tf32_error_model) and is therefore NOT the default: it is asserted at its own
documented bound of 1e-3 (the encoder-output bound) and offered as the fastest, lower-precision option.
Gradients as in tests/test_gpu_tf32.py / test_gpu_parity.py."""
import glob
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

import _golden
import _util
from polymer_chemprop_b200 import BatchMolGraph, MPN, synth

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)
MODEL_DIR = os.path.join(_golden.GOLDEN_DIR, "model")
NAMES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(MODEL_DIR, "*.npz")))


class Model(nn.Module):
    """MoleculeModel's regression forward (model.py:174-194): ffn(encoder(batch)); same parameter names."""

    def __init__(self, args, ffn_num_layers, ffn_hidden_size, num_tasks):
        super().__init__()
        self.encoder = MPN(args)                                        # model.py:41-47
        dropout, act = nn.Dropout(args.dropout), nn.ReLU()
        if ffn_num_layers == 1:                                         # model.py:82-102
            ffn = [dropout, nn.Linear(args.hidden_size, num_tasks)]
        else:
            ffn = [dropout, nn.Linear(args.hidden_size, ffn_hidden_size)]
            for _ in range(ffn_num_layers - 2):
                ffn += [act, dropout, nn.Linear(ffn_hidden_size, ffn_hidden_size)]
            ffn += [act, dropout, nn.Linear(ffn_hidden_size, num_tasks)]
        self.ffn = nn.Sequential(*ffn)

    def forward(self, batch):
        return self.ffn(self.encoder(batch))


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("mode,pred_tol,grad_tol", [("fp32", 1e-5, 1e-4), ("tf32x3", 1e-4, 5e-3), ("bf16x3", 1e-4, 5e-3), ("tf32", 1e-3, 3e-2)])
def test_full_model_predictions_and_gradients_match_reference(name, mode, pred_tol, grad_tol):
    """The default mode is the one that must meet the north star: MPN(args) without wd_math_mode."""
    z = np.load(os.path.join(MODEL_DIR, name + ".npz"))
    cfg = {str(k): _golden._parse(str(v)) for k, v in zip(z["cfg_keys"], z["cfg_vals"])}
    kind, n, seed = str(z["kind"]), int(z["n"]), int(z["seed"])
    bmg = BatchMolGraph(synth.make_graphs(kind, n, seed, random_edge_weights=(seed % 2 == 1)))
    assert (bmg.n_atoms, bmg.n_bonds) == (int(z["n_atoms"]), int(z["n_bonds"]))
    args = _util.make_args(hidden_size=cfg["hidden_size"], depth=cfg["depth"], bias=cfg.get("bias", False),
                           wd_math_mode=mode)
    model = Model(args, cfg["ffn_num_layers"], cfg["ffn_hidden_size"], cfg["num_tasks"])
    sd = {k[6:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("param:")}
    assert set(sd) == set(model.state_dict())                           # the reference's checkpoint keys, unchanged
    model.load_state_dict(sd)
    model.to(DEV).train()
    targets, mask = torch.from_numpy(z["targets"]).to(DEV), torch.from_numpy(z["mask"]).to(DEV)
    model.zero_grad()
    preds = model([bmg])
    loss = (nn.MSELoss(reduction="none")(preds, targets) * mask).sum() / mask.sum()     # train.py:70-71
    loss.backward()
    ref_preds = torch.from_numpy(z["preds"])
    _util.record("model_preds", mode=mode, name=name, rel=_golden.rel_err(preds.detach().cpu(), ref_preds),
                 max=_golden.max_err(preds.detach().cpu(), ref_preds))
    assert _golden.rel_err(preds.detach().cpu(), ref_preds) <= pred_tol
    assert _golden.max_err(preds.detach().cpu(), ref_preds) <= 20 * pred_tol
    assert abs(float(loss.detach()) - float(z["loss"])) <= 10 * pred_tol * abs(float(z["loss"]))
    print(f"{name} [{mode}]: prediction rel err {_golden.rel_err(preds.detach().cpu(), ref_preds):.2e}")
    grads = {k: p.grad.detach().cpu() for k, p in model.named_parameters() if p.grad is not None}
    ref_grads = {k[5:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("grad:")}
    assert set(grads) == set(ref_grads)
    _util.record("model_grads", mode=mode, name=name, grad=max(_golden.rel_err(grads[k], ref) for k, ref in ref_grads.items()))
    for k, ref in ref_grads.items():
        assert _golden.rel_err(grads[k], ref) <= grad_tol, k
    with torch.no_grad():                                               # predict.py:10-68 path (wd_encoder_infer)
        model.eval()
        assert _golden.rel_err(model([bmg]).cpu(), ref_preds) <= pred_tol


def test_default_math_mode_is_the_error_compensated_tensor_core_mode(monkeypatch):
    monkeypatch.delenv("WDMPNN_MATH", raising=False)
    args = _util.make_args()
    del args.wd_math_mode
    assert MPN(args).encoder[0].math_mode == "bf16x3"
