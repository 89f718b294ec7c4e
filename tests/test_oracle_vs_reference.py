"""CPU: the oracle restatement against the LIVE, unmodified reference module (imported through
oracle/ref_shim.py).  Skipped where /root/reference is absent (the GPU box); the committed
golden fixtures cover that case."""
import pytest
import torch

import _golden
from oracle import mpn_oracle, ref_shim
from polymer_chemprop_b200 import BatchMolGraph, synth

pytestmark = pytest.mark.skipif(not ref_shim.reference_available(), reason="reference tree not present")


@pytest.mark.parametrize("kind,n,cfg", [
    ("copolymer", 6, dict(hidden_size=48, depth=3)),
    ("copolymer", 5, dict(hidden_size=40, depth=4, undirected=True, bias=True, aggregation="sum")),
    ("qm9", 12, dict(hidden_size=32, depth=2, aggregation="norm", aggregation_norm=7)),
])
def test_oracle_equals_live_reference(kind, n, cfg):
    ref = ref_shim.load_reference()
    graphs = synth.make_graphs(kind, n, seed=11, dense_features=True, random_edge_weights=(kind == "copolymer"))
    ref_bmg = ref.BatchMolGraph(graphs)                 # the reference's own batch assembly
    ours_bmg = BatchMolGraph(graphs)                    # host mirror: identical tensors
    for x, y in zip(ref_bmg.get_components()[:7], ours_bmg.get_components()[:7]):
        assert torch.equal(x, y)
    assert list(ref_bmg.get_components()[7]) == list(ours_bmg.get_components()[7])
    torch.manual_seed(2)
    model = ref.MPN(ref_shim.make_args(**cfg), atom_fdim=ours_bmg.atom_fdim, bond_fdim=ours_bmg.bond_fdim)
    for p in model.parameters():
        if p.requires_grad:
            torch.nn.init.normal_(p, 0.0, 0.1) if p.dim() == 1 else torch.nn.init.xavier_normal_(p)
    out = model([ref_bmg])
    d_out = torch.randn(out.shape, generator=torch.Generator().manual_seed(4))
    out.backward(d_out)
    enc = model.encoder[0]
    weights = {k: v.detach() for k, v in enc.state_dict().items() if k != "cached_zero_vector"}
    o_out, o_grads = mpn_oracle.encoder_forward_backward(
        ours_bmg.get_components(), weights, d_out, depth=cfg["depth"], undirected=cfg.get("undirected", False),
        aggregation=cfg.get("aggregation", "mean"), aggregation_norm=cfg.get("aggregation_norm", 100))
    assert _golden.rel_err(o_out, out.detach()) <= 1e-6
    for k, p in enc.named_parameters():
        if p.grad is not None:
            assert _golden.rel_err(o_grads[k], p.grad) <= 1e-5, k
