"""GPU: wd_store_assemble (device-resident store -> one batch in HBM) against the host path for the same graphs in the
same order: every array bit-exact, then the encoder runs on it unchanged."""
import numpy as np
import pytest
import torch

import _golden
import _util
from polymer_chemprop_b200 import BatchMolGraph, DeviceGraph, DeviceGraphStore, synth

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)


def _graphs():
    return synth.make_graphs("copolymer", 400, 101, dense_features=True, random_edge_weights=True) + \
        [synth.empty_graph()] + synth.make_graphs("qm9", 100, 102, dense_features=True) + \
        [synth.single_atom_graph(np.random.default_rng(3))]


@pytest.mark.parametrize("n_sel,seed", [(1, 0), (37, 1), (600, 2)])
def test_assembled_batch_equals_host_built_batch_bit_for_bit(n_sel, seed):
    graphs = _graphs()
    store = DeviceGraphStore(graphs, DEV)
    idx = np.random.default_rng(seed).integers(0, len(graphs), size=n_sel)
    if n_sel > 10:
        idx[5] = 400                                                 # the empty molecule
    batch = store.batch(idx)
    got = DeviceGraph.get(batch, DEV)
    ref = DeviceGraph(BatchMolGraph([graphs[i] for i in idx]), DEV)
    ref.check()
    assert (got.n_atoms, got.n_bonds, got.n_mols) == (ref.n_atoms, ref.n_bonds, ref.n_mols)
    nb1 = max(ref.n_bonds - 1, 0)
    for name in ("f_atoms", "f_bonds", "w_atoms", "w_bonds", "in_ptr", "src", "rev", "dst", "mol_of", "a_start", "a_size", "dop"):
        a, b = getattr(got, name), getattr(ref, name)
        assert a.shape == b.shape and torch.equal(a.view(torch.int32), b.view(torch.int32)), name
    assert torch.equal(got.in_idx[:nb1], ref.in_idx[:nb1]) and torch.equal(got.out_idx[:nb1], ref.out_idx[:nb1])
    assert got.rev_adjacent == ref.rev_adjacent and got.struct.rev_adjacent == ref.struct.rev_adjacent
    assert batch.a_scope == BatchMolGraph([graphs[i] for i in idx]).a_scope


@pytest.mark.parametrize("mode", ["fp32", "tf32"])
def test_encoder_on_an_assembled_batch_equals_the_host_batch_bitwise(mode):
    graphs = _graphs()
    store = DeviceGraphStore(graphs, DEV)
    idx = np.random.default_rng(7).permutation(len(graphs))[:300]
    torch.manual_seed(0)
    model = _util.MPN(_util.make_args(hidden_size=300, depth=3, wd_math_mode=mode), atom_fdim=store.atom_fdim,
                      bond_fdim=store.bond_fdim).to(DEV)
    for p in model.parameters():
        if p.requires_grad and p.dim() == 2:
            torch.nn.init.xavier_normal_(p)
    d_out = torch.randn(300, 300, generator=torch.Generator().manual_seed(1))
    side = torch.cuda.Stream(device=DEV)
    out_s, grads_s = _util.run_fwd_bwd(model, store.batch(idx, stream=side), d_out)     # assembled on a prefetch stream
    out_h, grads_h = _util.run_fwd_bwd(model, BatchMolGraph([graphs[i] for i in idx]), d_out)
    assert torch.equal(out_s, out_h)
    for k in grads_h:
        assert torch.equal(grads_s[k], grads_h[k]), k


def test_repeated_store_assembles_the_same_batches_as_a_store_of_the_copies():
    """DeviceGraphStore(graphs, repeat=k) (how bench.py stages a dataset of benchmark size) == a store packed from the
    k copies one by one: same arrays on the host, and a batch drawn across the copies assembles bit-identically."""
    graphs = _graphs()[:150]
    tiled = DeviceGraphStore._tile(DeviceGraphStore.pack(graphs), 3)
    full = DeviceGraphStore.pack(graphs * 3)
    for k, v in full.items():
        if isinstance(v, np.ndarray):
            assert np.array_equal(tiled[k], v), k
        else:
            assert tiled[k] == v, k
    store = DeviceGraphStore(graphs, DEV, repeat=3)
    assert len(store) == 450
    idx = np.random.default_rng(5).integers(0, 450, size=200)
    got = DeviceGraph.get(store.batch(idx), DEV)
    ref = DeviceGraph(BatchMolGraph([graphs[i % 150] for i in idx]), DEV)
    for name in ("f_atoms", "f_bonds", "w_bonds", "in_ptr", "src", "rev", "dop"):
        assert torch.equal(getattr(got, name).view(torch.int32), getattr(ref, name).view(torch.int32)), name
