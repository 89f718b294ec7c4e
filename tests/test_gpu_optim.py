"""GPU parity of the fused optimiser step (wd_adam_step through the C ABI, FlatAdam) with the CPU oracle and with
torch.optim.Adam / AdamW + clip_grad_norm_ -- what chemprop/utils.py:295-310 and chemprop/train/train.py:80-86 run.
Tolerance: 2e-6 relative to max(1, |p|) per step (fp32; the fused kernel uses FMAs where torch rounds twice)."""
import ctypes as C

import numpy as np
import pytest
import torch

import _util
from oracle import optim_oracle
from polymer_chemprop_b200 import BatchMolGraph, FlatAdam, NoamLR, _capi, dp, synth

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)
TOL = 2e-6


def _step(p, g, m, v, step, scratch, norm, **kw):
    cfg = _capi.WdAdam(kw.get("lr", 1e-3), kw.get("beta1", 0.9), kw.get("beta2", 0.999), kw.get("eps", 1e-8),
                       kw.get("weight_decay", 0.0), kw.get("max_grad_norm") or 0.0, step, int(kw.get("decoupled", False)))
    _capi.check(_capi.lib().wd_adam_step(_capi.ptr(p), _capi.ptr(g), _capi.ptr(m), _capi.ptr(v), p.numel(), C.byref(cfg),
                                         _capi.ptr(scratch), _capi.ptr(norm), _capi.current_stream(DEV)))


@pytest.mark.parametrize("n", [1, 3, 4, 1027, 271_500])
@pytest.mark.parametrize("kw", [dict(), dict(weight_decay=0.01), dict(weight_decay=0.05, decoupled=True, lr=3e-3),
                                dict(max_grad_norm=0.7), dict(max_grad_norm=1e9, weight_decay=0.01)])
def test_adam_step_matches_oracle(n, kw):
    rng = np.random.default_rng(n)
    p0 = rng.standard_normal(n).astype(np.float32)
    p, m, v = (torch.from_numpy(x).to(DEV) for x in (p0, np.zeros(n, np.float32), np.zeros(n, np.float32)))
    scratch = torch.zeros(int(_capi.lib().wd_adam_scratch_floats()), device=DEV)
    norm = torch.zeros(1, device=DEV)
    rp, rm, rv = p0, np.zeros(n, np.float32), np.zeros(n, np.float32)
    for step in range(1, 9):
        g = (rng.standard_normal(n) * (3.0 if step == 4 else 1.0)).astype(np.float32)
        _step(p, torch.from_numpy(g).to(DEV), m, v, step, scratch, norm, **kw)
        rp, rm, rv = optim_oracle.adam_step(rp, g, rm, rv, step, **kw)
        if kw.get("max_grad_norm"):
            assert norm.item() == pytest.approx(float(np.sqrt(np.sum(g.astype(np.float64) ** 2))), rel=1e-5)
        assert np.max(np.abs(p.cpu().numpy() - rp)) <= TOL * max(1.0, np.max(np.abs(rp))), step
        assert np.max(np.abs(m.cpu().numpy() - rm)) <= TOL * max(1.0, np.max(np.abs(rm))), step
        assert np.max(np.abs(v.cpu().numpy() - rv)) <= TOL * max(1.0, np.max(np.abs(rv))), step


def test_adam_step_is_bitwise_repeatable_and_validates_its_arguments():
    n = 50_001
    g = torch.randn(n, device=DEV, generator=torch.Generator(device=DEV).manual_seed(1))
    outs = []
    for _ in range(2):
        p, m, v = torch.ones(n, device=DEV), torch.zeros(n, device=DEV), torch.zeros(n, device=DEV)
        scratch, norm = torch.zeros(int(_capi.lib().wd_adam_scratch_floats()), device=DEV), torch.zeros(1, device=DEV)
        for step in (1, 2, 3):
            _step(p, g, m, v, step, scratch, norm, max_grad_norm=0.3)
        outs.append((p.clone(), norm.clone()))
    assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    cfg = _capi.WdAdam(1e-3, 0.9, 0.999, 1e-8, 0.0, 0.0, 0, 0)                      # step 0
    rc = _capi.lib().wd_adam_step(_capi.ptr(p), _capi.ptr(g), _capi.ptr(m), _capi.ptr(v), n, C.byref(cfg), None, None,
                                  _capi.current_stream(DEV))
    assert rc != 0 and b"step" in _capi.lib().wd_last_error()


def _models(hidden=64, depth=3):
    graphs = synth.make_graphs("copolymer", 24, 11, dense_features=True, random_edge_weights=True)
    bmg = BatchMolGraph(graphs)
    torch.manual_seed(4)
    a = _util.MPN(_util.make_args(hidden_size=hidden, depth=depth, bias=True), atom_fdim=bmg.atom_fdim,
                  bond_fdim=bmg.bond_fdim).to(DEV)
    b = _util.MPN(_util.make_args(hidden_size=hidden, depth=depth, bias=True), atom_fdim=bmg.atom_fdim,
                  bond_fdim=bmg.bond_fdim).to(DEV)
    b.load_state_dict(a.state_dict())
    return bmg, a, b


@pytest.mark.parametrize("kind,wd,clip", [("adam", 0.0, None), ("adamw", 0.01, 0.05)])
def test_flat_adam_trains_the_encoder_like_torch_adam_under_noam(kind, wd, clip):
    """The reference's loop (train.py:56-86): zero_grad, forward, backward, clip, optimizer.step, scheduler.step --
    once with torch's optimiser, once with FlatAdam + the NoamLR mirror, same model, same batches."""
    bmg, ref, ours = _models()
    params_ref = [p for p in ref.parameters() if p.requires_grad]
    params = [p for p in ours.parameters() if p.requires_grad]
    cls = torch.optim.AdamW if kind == "adamw" else torch.optim.Adam
    opt_ref = cls([{"params": params_ref, "lr": 1e-4, "weight_decay": wd}])
    opt = FlatAdam([{"params": params, "lr": 1e-4, "weight_decay": wd}], decoupled=kind == "adamw", max_grad_norm=clip)
    sch_ref = NoamLR(opt_ref, [1.0], [4], 3, [1e-4], [1e-3], [1e-4])
    sch = NoamLR(opt, [1.0], [4], 3, [1e-4], [1e-3], [1e-4])
    assert all(p.data_ptr() >= opt.flat_params.data_ptr() for p in params)           # the module reads the flat buffer
    gen = torch.Generator().manual_seed(8)
    for step in range(10):
        d_out = torch.randn(len(bmg.a_scope), 64, generator=gen).to(DEV)
        for model, o, s, ps in ((ref, opt_ref, sch_ref, params_ref), (ours, opt, sch, params)):
            model.zero_grad()
            out = model([bmg])
            out.backward(d_out)
            if clip and o is opt_ref:
                torch.nn.utils.clip_grad_norm_(ps, clip)
            o.step()
            s.step()
        assert opt.param_groups[0]["lr"] == opt_ref.param_groups[0]["lr"]
        for (k, a), b in zip(ref.named_parameters(), ours.parameters()):
            assert torch.max(torch.abs(a - b)).item() <= 2e-5 * max(1.0, a.abs().max().item()), (step, k)
    sd = opt.state_dict()
    assert len(sd["state"]) == len(params) and sd["param_groups"][0]["lr"] == opt.param_groups[0]["lr"]


def test_flat_adam_reads_the_all_reduce_buffer_directly():
    bmg, ref, ours = _models(hidden=32, depth=2)
    params = [p for p in ours.parameters() if p.requires_grad]
    reducer = dp.FlatGradAllReduce(params)
    opt = FlatAdam(params, lr=1e-3, grad_buffer=reducer)
    assert opt.flat_grads.data_ptr() == reducer.flat.data_ptr()
    opt_ref = torch.optim.Adam([p for p in ref.parameters() if p.requires_grad], lr=1e-3)
    d_out = torch.randn(len(bmg.a_scope), 32, generator=torch.Generator().manual_seed(2)).to(DEV)
    for _ in range(3):
        for model, o in ((ref, opt_ref), (ours, opt)):
            model.zero_grad()
            model([bmg]).backward(d_out)
            if o is opt:
                reducer()
            o.step()
    for a, b in zip(ref.parameters(), ours.parameters()):
        assert torch.max(torch.abs(a - b)).item() <= 2e-5 * max(1.0, a.abs().max().item())
