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


class _MaskedDropout(torch.nn.Module):
    """Stand-in for the reference's nn.Dropout whose k-th call multiplies by the k-th given mask: shares the
    product's Philox masks with the UNMODIFIED reference forward so that the oracle's dropout placement
    (mpn.py:124 after every message update, :134 after the atom update) is pinned against it."""

    def __init__(self, multipliers):
        super().__init__()
        self.multipliers, self.calls = multipliers, 0

    def forward(self, x):
        m = self.multipliers[self.calls]
        self.calls += 1
        return x * m


def test_oracle_dropout_placement_equals_live_reference():
    ref = ref_shim.load_reference()
    graphs = synth.make_graphs("copolymer", 6, seed=3, dense_features=True, random_edge_weights=True)
    ref_bmg, ours_bmg = ref.BatchMolGraph(graphs), BatchMolGraph(graphs)
    h, depth, p, lo, hi = 48, 4, 0.35, 0x1234ABCD, 0x0BADF00D
    torch.manual_seed(6)
    model = ref.MPN(ref_shim.make_args(hidden_size=h, depth=depth, dropout=p), atom_fdim=ours_bmg.atom_fdim,
                    bond_fdim=ours_bmg.bond_fdim)
    for q in model.parameters():
        if q.requires_grad:
            torch.nn.init.normal_(q, 0.0, 0.1) if q.dim() == 1 else torch.nn.init.xavier_normal_(q)
    enc = model.encoder[0]
    ld = (h + 7) // 8 * 8
    n_b, n_a = ours_bmg.n_bonds, ours_bmg.n_atoms
    masks = [mpn_oracle.dropout_multiplier(n_b, h, ld, p, lo, hi, t) for t in range(1, depth)] + \
            [mpn_oracle.dropout_multiplier(n_a, h, ld, p, lo, hi, depth)]
    assert 0.25 < float((masks[0] == 0).float().mean()) < 0.45
    enc.dropout_layer = _MaskedDropout(masks)
    model.train()
    out = model([ref_bmg])
    assert enc.dropout_layer.calls == depth
    d_out = torch.randn(out.shape, generator=torch.Generator().manual_seed(4))
    out.backward(d_out)
    weights = {k: v.detach() for k, v in enc.state_dict().items() if k != "cached_zero_vector"}
    o_out, o_grads = mpn_oracle.encoder_forward_backward(
        ours_bmg.get_components(), weights, d_out, depth=depth,
        dropout=dict(p=p, seed_lo=lo, seed_hi=hi, ld=ld))
    assert _golden.rel_err(o_out, out.detach()) <= 1e-6
    for k, q in enc.named_parameters():
        if q.grad is not None:
            assert _golden.rel_err(o_grads[k], q.grad) <= 1e-5, k


def test_install_makes_the_reference_model_build_this_encoder():
    """mpn.install(): chemprop.models.model.MoleculeModel.create_encoder (model.py:41-47) then
    instantiates this package's MPN; the reference's checkpoint keys / shapes are unchanged, so its own
    load_checkpoint (utils.py:98-135) keeps working.  Host logic only -- no GPU compute here."""
    import importlib
    from oracle import make_golden_model
    from polymer_chemprop_b200 import mpn as ours
    ref_shim.load_reference()
    cm = importlib.import_module("chemprop.models.model")
    cmpn = importlib.import_module("chemprop.models.mpn")
    args = make_golden_model.model_args(hidden_size=64, depth=3, ffn_num_layers=2, ffn_hidden_size=32, num_tasks=2)
    torch.manual_seed(0)
    ref_model = cm.MoleculeModel(args)
    saved = (cm.MPN, cmpn.MPN, cmpn.MPNEncoder)
    try:
        ours.install()
        model = cm.MoleculeModel(args)
        assert isinstance(model.encoder, ours.MPN) and isinstance(model.encoder.encoder[0], ours.MPNEncoder)
        assert list(model.state_dict()) == list(ref_model.state_dict())
        assert all(model.state_dict()[k].shape == v.shape for k, v in ref_model.state_dict().items())
        model.load_state_dict(ref_model.state_dict())
        with pytest.raises(RuntimeError, match="no CPU fallback"):      # parameters on the CPU: refuse, never fall back
            model([BatchMolGraph(synth.make_graphs("qm9", 2, 0))])
    finally:
        cm.MPN, cmpn.MPN, cmpn.MPNEncoder = saved


def test_install_swaps_batch_mol_graph_and_it_takes_reference_mol_graphs():
    """install() also puts this package's BatchMolGraph where the reference's loader builds batches
    (chemprop/data/data.py:312 resolves the name in chemprop.data.data; mol2graph in features.featurization).  Built
    from the reference's own MolGraph-style objects (Python lists), it must equal the reference's BatchMolGraph
    attribute for attribute -- and offer the compact upload form on top."""
    import importlib
    from polymer_chemprop_b200 import mpn as ours
    ref_shim.load_reference()
    feat = importlib.import_module("chemprop.features.featurization")
    mods = [importlib.import_module(n) for n in ("chemprop.features.featurization", "chemprop.features", "chemprop.data.data",
                                                 "chemprop.models.model", "chemprop.models.mpn")]
    RefBMG = feat.BatchMolGraph
    graphs = synth.make_graphs("copolymer", 9, 4, random_edge_weights=True) + [synth.empty_graph()] + \
        synth.make_graphs("qm9", 3, 5)
    ref_style = [g.as_reference_lists() for g in graphs]          # what MolGraph.__init__ leaves behind: lists
    ref_bmg = RefBMG(ref_style)
    saved = [(m, m.BatchMolGraph) for m in mods if hasattr(m, "BatchMolGraph")]
    saved_mpn = (mods[3].MPN, mods[4].MPN, mods[4].MPNEncoder)
    try:
        ours.install()
        assert all(m.BatchMolGraph is BatchMolGraph for m, _ in saved) and len(saved) >= 3
        new_bmg = mods[2].BatchMolGraph(ref_style)                # the name data.py:312 resolves
        a, b = ref_bmg.get_components(), new_bmg.get_components()
        for x, y in zip(a[:7], b[:7]):
            assert x.dtype == y.dtype and torch.equal(x, y)
        assert list(a[7]) == list(b[7]) and list(a[8]) == list(b[8])
        assert [float(v) for v in a[9]] == [float(v) for v in b[9]]
        assert (new_bmg.n_atoms, new_bmg.n_bonds, new_bmg.max_num_bonds) == (ref_bmg.n_atoms, ref_bmg.n_bonds, ref_bmg.max_num_bonds)
        assert new_bmg.f_bonds_tail is not None and new_bmg.a_scope_array is not None      # the compact upload form
        assert all(hasattr(g, "_wd_data") for g in ref_style)     # converted once per molecule, cached on it
    finally:
        for m, v in saved:
            m.BatchMolGraph = v
        mods[3].MPN, mods[4].MPN, mods[4].MPNEncoder = saved_mpn


def test_noam_lr_oracle_and_mirror_equal_live_reference_scheduler():
    """chemprop/nn_utils.py:115-194 driven as chemprop/train/train.py:82-86 drives it, next to the product's mirror
    and the oracle formula: equal learning rates to the last bit at every step, including the step the base class
    takes inside the constructor."""
    import importlib
    from oracle import optim_oracle
    from polymer_chemprop_b200 import NoamLR
    ref_shim.load_reference()
    RefNoam = importlib.import_module("chemprop.nn_utils").NoamLR
    def cfg():   # fresh lists per scheduler: the reference keeps (and then overwrites) the caller's init_lr list (:157, :188)
        return dict(warmup_epochs=[1.5], total_epochs=[5], steps_per_epoch=11, init_lr=[2e-4], max_lr=[3e-3], final_lr=[5e-5])
    opts = [torch.optim.Adam([torch.nn.Parameter(torch.zeros(2))], lr=2e-4) for _ in range(2)]
    a, b = RefNoam(optimizer=opts[0], **cfg()), NoamLR(opts[1], **cfg())
    for _ in range(60):
        assert a.current_step == b.current_step
        assert opts[0].param_groups[0]["lr"] == opts[1].param_groups[0]["lr"] == a.get_lr()[0] == b.get_lr()[0]
        assert optim_oracle.noam_lr(a.current_step, 1.5, 5, 11, 2e-4, 3e-3, 5e-5) == opts[0].param_groups[0]["lr"]
        for o in opts:
            o.param_groups[0]["params"][0].grad = torch.ones(2)
            o.step()
        a.step()
        b.step()
