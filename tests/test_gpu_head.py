"""GPU: the fused FFN head + masked MSE loss (wd_head_forward / wd_head_backward, polymer_chemprop_b200/head.py) against
(1) torch's own float64 autograd of the reference's expressions (model.py:174-194, train.py:70-74) on random data, and
(2) the UNMODIFIED reference MoleculeModel's golden predictions, loss and parameter gradients (tests/golden/model/),
with the encoder in front of it.  Tolerances: the head is fp32 CUDA-core arithmetic -> 1e-5 relative on predictions /
loss and 1e-4 on gradients (the FP32 bounds of tests/test_gpu_parity.py); behind the tf32x3 encoder the bounds of
tests/test_gpu_model.py apply (predictions 1e-4 = the north star's bound, gradients 5e-3)."""
import glob
import os

import numpy as np
import pytest
import torch
import torch.nn as nn

import _golden
import _util
from polymer_chemprop_b200 import BatchMolGraph, MPN, ffn_mse_loss, ffn_predict, synth

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)
MODEL_DIR = os.path.join(_golden.GOLDEN_DIR, "model")
NAMES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(MODEL_DIR, "*.npz")))
ACTS = {"ReLU": nn.ReLU, "LeakyReLU": lambda: nn.LeakyReLU(0.1), "tanh": nn.Tanh, "SELU": nn.SELU, "ELU": nn.ELU}


def _ffn(in_dim, hidden, out_dim, n_layers, act):
    """create_ffn's layout (model.py:82-102)."""
    drop = nn.Dropout(0.0)
    if n_layers == 1:
        mods = [drop, nn.Linear(in_dim, out_dim)]
    else:
        mods = [drop, nn.Linear(in_dim, hidden)]
        for _ in range(n_layers - 2):
            mods += [ACTS[act](), drop, nn.Linear(hidden, hidden)]
        mods += [ACTS[act](), drop, nn.Linear(hidden, out_dim)]
    return nn.Sequential(*mods)


@pytest.mark.parametrize("G,in_dim,hidden,T,n_layers,act", [(50, 300, 300, 1, 2, "ReLU"), (4096, 300, 300, 1, 2, "ReLU"),
                                                          (37, 64, 48, 3, 3, "tanh"), (129, 76, 20, 2, 1, "ReLU"),
                                                          (700, 600, 300, 12, 4, "ELU"), (3, 8, 5, 1, 2, "LeakyReLU"),
                                                          (300, 100, 33, 2, 2, "SELU")])
def test_head_matches_float64_autograd(G, in_dim, hidden, T, n_layers, act):
    torch.manual_seed(G + T)
    ffn = _ffn(in_dim, hidden, T, n_layers, act).to(DEV)
    gen = torch.Generator().manual_seed(1)
    x = torch.randn(G, in_dim, generator=gen)
    targets = torch.randn(G, T, generator=gen)
    mask = torch.rand(G, T, generator=gen) > 0.2
    mask[0, 0] = True
    tw, dw = torch.rand(T, generator=gen) + 0.5, torch.rand(G, 1, generator=gen) + 0.5
    # float64 reference: the reference's own expressions
    ref_ffn = _ffn(in_dim, hidden, T, n_layers, act).double()
    ref_ffn.load_state_dict({k: v.double().cpu() for k, v in ffn.state_dict().items()})
    xr = x.double().requires_grad_(True)
    pr = ref_ffn(xr)
    lr = (nn.MSELoss(reduction="none")(pr, targets.double()) * tw.double() * dw.double() * mask).sum() / mask.sum()
    lr.backward()
    xg = x.to(DEV).requires_grad_(True)
    preds, loss = ffn_mse_loss(xg, ffn, targets, mask, target_weights=tw, data_weights=dw)
    loss.backward()
    assert _golden.rel_err(preds.detach().cpu(), pr.detach()) <= 1e-5
    assert abs(float(loss.detach()) - float(lr.detach())) <= 1e-5 * abs(float(lr.detach()))
    assert _golden.rel_err(xg.grad.cpu(), xr.grad) <= 1e-4
    for (k, p), (_k, q) in zip(ffn.named_parameters(), ref_ffn.named_parameters()):
        assert _golden.rel_err(p.grad.cpu(), q.grad) <= 1e-4, k
    assert _golden.rel_err(ffn_predict(xg, ffn).cpu(), pr.detach()) <= 1e-5
    # gradient flowing in through preds as well as through the loss (both outputs are differentiable)
    for p in ffn.parameters():
        p.grad = None
    preds2, loss2 = ffn_mse_loss(x.to(DEV), ffn, targets, mask)
    (3.0 * loss2 + preds2.sum()).backward()
    for q in ref_ffn.parameters():
        q.grad = None
    pr2 = ref_ffn(x.double())
    (3.0 * (nn.MSELoss(reduction="none")(pr2, targets.double()) * mask).sum() / mask.sum() + pr2.sum()).backward()
    for (k, p), (_k, q) in zip(ffn.named_parameters(), ref_ffn.named_parameters()):
        assert _golden.rel_err(p.grad.cpu(), q.grad) <= 1e-4, k


def test_head_is_bitwise_deterministic_and_refuses_what_it_does_not_run():
    ffn = _ffn(300, 300, 1, 2, "ReLU").to(DEV)
    x = torch.randn(500, 300, device=DEV)
    t, m = torch.randn(500, 1), torch.ones(500, 1, dtype=torch.bool)
    runs = []
    for _ in range(2):
        ffn.zero_grad()
        xg = x.clone().requires_grad_(True)
        p, l = ffn_mse_loss(xg, ffn, t, m)
        l.backward()
        runs.append([p.detach().clone(), l.detach().clone(), xg.grad.clone()] + [q.grad.clone() for q in ffn.parameters()])
    assert all(torch.equal(a, b) for a, b in zip(*runs))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ffn_mse_loss(x.cpu(), ffn, t, m)
    bad = nn.Sequential(nn.Dropout(0.3), nn.Linear(300, 1)).to(DEV).train()
    with pytest.raises(NotImplementedError, match="dropout"):
        ffn_mse_loss(x, bad, t, m)
    with pytest.raises(NotImplementedError):
        ffn_mse_loss(x, nn.Sequential(nn.Linear(300, 1), nn.Sigmoid()).to(DEV), t, m)


@pytest.mark.parametrize("name", NAMES)
@pytest.mark.parametrize("mode,pred_tol,grad_tol", [("fp32", 1e-5, 1e-4), ("tf32x3", 1e-4, 5e-3)])
def test_encoder_plus_fused_head_matches_the_reference_models_golden_step(name, mode, pred_tol, grad_tol):
    """One regression training step of the UNMODIFIED reference MoleculeModel (oracle/make_golden_model.py): this
    package's MPN + the reference's own nn.Sequential FFN run by the fused head -> predictions, loss, every gradient."""
    from test_gpu_model import Model
    z = np.load(os.path.join(MODEL_DIR, name + ".npz"))
    cfg = {str(k): _golden._parse(str(v)) for k, v in zip(z["cfg_keys"], z["cfg_vals"])}
    kind, n, seed = str(z["kind"]), int(z["n"]), int(z["seed"])
    bmg = BatchMolGraph(synth.make_graphs(kind, n, seed, random_edge_weights=(seed % 2 == 1)))
    args = _util.make_args(hidden_size=cfg["hidden_size"], depth=cfg["depth"], bias=cfg.get("bias", False), wd_math_mode=mode)
    model = Model(args, cfg["ffn_num_layers"], cfg["ffn_hidden_size"], cfg["num_tasks"])
    model.load_state_dict({k[6:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("param:")})
    model.to(DEV).train()
    targets, mask = torch.from_numpy(z["targets"]), torch.from_numpy(z["mask"])
    model.zero_grad()
    preds, loss = ffn_mse_loss(model.encoder([bmg]), model.ffn, targets, mask)
    loss.backward()
    ref_preds = torch.from_numpy(z["preds"])
    _util.record("fused_head_model", mode=mode, name=name, preds=_golden.rel_err(preds.detach().cpu(), ref_preds))
    assert _golden.rel_err(preds.detach().cpu(), ref_preds) <= pred_tol
    assert abs(float(loss.detach()) - float(z["loss"])) <= 10 * pred_tol * abs(float(z["loss"]))
    grads = {k: p.grad.detach().cpu() for k, p in model.named_parameters() if p.grad is not None}
    ref_grads = {k[5:]: torch.from_numpy(z[k]) for k in z.files if k.startswith("grad:")}
    assert set(grads) == set(ref_grads)
    for k, ref in ref_grads.items():
        assert _golden.rel_err(grads[k], ref) <= grad_tol, k
