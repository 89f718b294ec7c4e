"""Shared helpers for the GPU parity tests."""
import types

import torch

from polymer_chemprop_b200 import BatchMolGraph, MPN


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
