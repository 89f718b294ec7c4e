"""Shared helpers for the GPU parity tests."""
import json
import os
import types

import torch

import _golden
from polymer_chemprop_b200 import BatchMolGraph, MPN, synth


def make_args(hidden_size=300, depth=3, bias=False, dropout=0.0, activation="ReLU", undirected=False,
              aggregation="mean", aggregation_norm=100, device="cuda", number_of_molecules=1,
              mpn_shared=False, wd_math_mode="fp32"):
    """The TrainArgs fields the encoder reads (chemprop/models/mpn.py:26-35, :189-208)."""
    return types.SimpleNamespace(
        hidden_size=hidden_size, depth=depth, bias=bias, dropout=dropout, activation=activation,
        atom_messages=False, undirected=undirected, aggregation=aggregation,
        aggregation_norm=aggregation_norm, atom_descriptors=None, atom_descriptors_size=0,
        device=torch.device(device), mpn_shared=mpn_shared, number_of_molecules=number_of_molecules,
        features_only=False, use_input_features=False, overwrite_default_atom_features=False,
        overwrite_default_bond_features=False, wd_math_mode=wd_math_mode)


class ComponentsGraph:
    """Duck-typed BatchMolGraph built from a get_components() tuple (golden fixtures)."""

    def __init__(self, components):
        self._c = components

    def get_components(self, atom_messages=False):
        return self._c


def model_from_params(params, cfg, math_mode="fp32", device="cuda"):
    args = make_args(device=device, wd_math_mode=math_mode, **cfg)
    fa = params["W_o.weight"].shape[1] - params["W_o.weight"].shape[0]
    fb = params["W_i.weight"].shape[1]
    model = MPN(args, atom_fdim=fa, bond_fdim=fb)
    model.encoder[0].load_state_dict(params)
    return model.to(device)


def run_fwd_bwd(model, graph, d_out):
    model.zero_grad()
    out = model([graph])
    out.backward(d_out.to(out.device))
    grads = {k: p.grad.detach().cpu() for k, p in model.encoder[0].named_parameters() if p.grad is not None}
    return out.detach().cpu(), grads


def dropout_case(math_mode, hidden, depth, activation, p, out_tol, grad_tol):
    """Training-mode dropout against the oracle run with the same Philox masks (GPU tests only)."""
    from oracle import mpn_oracle
    graphs = synth.make_graphs("copolymer", 40, 31, dense_features=True, random_edge_weights=True)
    bmg = BatchMolGraph(graphs + [synth.empty_graph()])
    torch.manual_seed(5)
    model = MPN(make_args(hidden_size=hidden, depth=depth, activation=activation, dropout=p,
                                      wd_math_mode=math_mode), atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(torch.device('cuda', 0))
    for q in model.parameters():
        if q.requires_grad:
            torch.nn.init.normal_(q, 0.0, 0.1) if q.dim() == 1 else torch.nn.init.xavier_normal_(q)
    d_out = torch.randn(len(bmg.a_scope), hidden, generator=torch.Generator().manual_seed(9))
    enc = model.encoder[0]
    weights = {k: v.detach().cpu() for k, v in enc.state_dict().items() if k != "cached_zero_vector"}
    model.train()
    out, grads = run_fwd_bwd(model, bmg, d_out)
    cfg = enc.last_config
    assert abs(cfg.dropout - p) < 1e-7
    drop = dict(p=p, seed_lo=cfg.dropout_seed_lo, seed_hi=cfg.dropout_seed_hi, ld=(hidden + 7) // 8 * 8)
    ref_out, ref_grads = mpn_oracle.encoder_forward_backward(bmg.get_components(), weights, d_out, depth=depth,
                                                             activation=activation, dropout=drop)
    assert _golden.rel_err(out, ref_out) <= out_tol
    for k, ref in ref_grads.items():
        assert _golden.rel_err(grads[k], ref) <= grad_tol, k
    # a second training call draws a new key; eval() switches dropout off (nn.Dropout semantics, mpn.py:38)
    out2, _ = run_fwd_bwd(model, bmg, d_out)
    assert (enc.last_config.dropout_seed_lo, enc.last_config.dropout_seed_hi) != (drop["seed_lo"], drop["seed_hi"])
    assert not torch.equal(out, out2)
    model.eval()
    out_eval, grads_eval = run_fwd_bwd(model, bmg, d_out)
    ref_eval, ref_grads_eval = mpn_oracle.encoder_forward_backward(bmg.get_components(), weights, d_out, depth=depth,
                                                                   activation=activation)
    assert enc.last_config.dropout == 0.0 and _golden.rel_err(out_eval, ref_eval) <= out_tol
    # the same torch seed reproduces the same masks
    model.train()
    torch.manual_seed(77)
    a, _ = run_fwd_bwd(model, bmg, d_out)
    torch.manual_seed(77)
    b, _ = run_fwd_bwd(model, bmg, d_out)
    assert torch.equal(a, b)


# tolerances per math mode: (encoder output, gradients on the tiny golden batches, raw contraction) -- relative L2.
# tf32x3 (the default mode): error-compensated forward, fp32-grade output; its backward contractions are single-pass
# TF32 on exact activations (no ReLU-mask flips), hence the tighter gradient bound than plain tf32.
TOL = {"fp32": dict(out=1e-5, grad=1e-4, gemm=1e-5),
       "tf32": dict(out=1e-3, grad=3e-2, gemm=1e-3),
       "tf32x3": dict(out=1e-5, grad=5e-3, gemm=5e-6),
       # bf16x3: the same three products with the split point at bfloat16 (2^-17 dropped instead of 2^-22): the
       # exact-accumulation model gives 3e-6 on the encoder output (tests/test_oracle_golden.py)
       "bf16x3": dict(out=2e-5, grad=5e-3, gemm=2e-5)}


def record(key, **values):
    """Appends measured errors to gpurun_out/test_metrics.jsonl (when that directory exists) so that the tolerances
    written in the tests can be compared with what the B200 run actually measured."""
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    if os.path.isdir(root):
        with open(os.path.join(root, "test_metrics.jsonl"), "a") as f:
            f.write(json.dumps({"key": key, **values}) + "\n")
