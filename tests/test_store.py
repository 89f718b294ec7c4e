"""CPU: the compact packing behind DeviceGraphStore and the assembly formulae of csrc/store.cu, restated in numpy,
reproduce what BatchMolGraph (pinned bit-exactly to the reference's, tests/test_batch_graph.py) + the CSR oracle
give for the same graphs in the same order."""
import numpy as np
import pytest
import torch

from oracle import mpn_oracle
from polymer_chemprop_b200 import BatchMolGraph, synth
from polymer_chemprop_b200.store import DeviceGraphStore


def assemble_numpy(pk, idx):
    """csrc/store.cu (k_store_atoms / k_store_bonds / k_store_mols) in numpy."""
    na, nb = pk["n_atoms_of"][idx], pk["n_bonds_of"][idx]
    a_start = 1 + np.concatenate([[0], np.cumsum(na)[:-1]])
    b_start = 1 + np.concatenate([[0], np.cumsum(nb)[:-1]])
    Na, Nb, Fa, Ft = int(1 + na.sum()), int(1 + nb.sum()), pk["atom_fdim"], pk["tail_fdim"]
    out = dict(f_atoms=np.zeros((Na, Fa), np.float32), f_bonds=np.zeros((Nb, Fa + Ft), np.float32),
               w_atoms=np.zeros(Na, np.float32), w_bonds=np.zeros(Nb, np.float32), in_ptr=np.zeros(Na + 1, np.int32),
               in_idx=np.zeros(max(Nb - 1, 1), np.int32), out_idx=np.zeros(max(Nb - 1, 1), np.int32),
               src=np.zeros(Nb, np.int32), rev=np.zeros(Nb, np.int32), dst=np.zeros(Nb, np.int32),
               mol_of=np.full(Na, -1, np.int32), dop=pk["dop"][idx])
    out["in_ptr"][Na] = Nb - 1
    for g, s in enumerate(idx):
        a0, b0 = pk["atom_off"][s], pk["bond_off"][s]
        for la in range(na[g]):
            r = a_start[g] + la
            out["f_atoms"][r] = pk["f_atoms"][a0 + la]
            out["w_atoms"][r] = pk["w_atoms"][a0 + la]
            out["in_ptr"][r] = b_start[g] - 1 + pk["in_ptr_local"][a0 + la]
            out["mol_of"][r] = g
        for lb in range(nb[g]):
            b, sb = b_start[g] + lb, b0 + lb
            la, lr = pk["b2a"][sb], pk["b2revb"][sb]
            out["f_bonds"][b] = np.concatenate([pk["f_atoms"][a0 + la], pk["bond_tail"][sb, :Ft]])
            out["w_bonds"][b] = pk["w_bonds"][sb]
            out["src"][b], out["rev"][b] = a_start[g] + la, b_start[g] + lr
            out["dst"][b] = a_start[g] + pk["b2a"][b0 + lr]
            out["in_idx"][b - 1] = b_start[g] + pk["in_idx"][sb]
            out["out_idx"][b - 1] = b_start[g] + pk["out_idx"][sb]
    return out, a_start, na


@pytest.mark.parametrize("seed", [0, 1])
def test_packed_store_assembles_to_the_reference_batch(seed):
    graphs = synth.make_graphs("copolymer", 12, 90 + seed, dense_features=True, random_edge_weights=True) + \
        [synth.empty_graph()] + synth.make_graphs("qm9", 6, 95 + seed, dense_features=True) + \
        [synth.single_atom_graph(np.random.default_rng(3))]
    pk = DeviceGraphStore.pack(graphs)
    assert pk["rev_adjacent"]
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, len(graphs), size=15)                      # repeats and any order
    idx[3] = 12                                                      # the empty molecule
    got, a_start, na = assemble_numpy(pk, idx)
    bmg = BatchMolGraph([graphs[i] for i in idx])
    fa, fb, wa, wb, a2b, b2a, b2revb, a_scope, _b_scope, dop = bmg.get_components()
    assert np.array_equal(got["f_atoms"].view(np.uint32), fa.numpy().view(np.uint32))
    assert np.array_equal(got["f_bonds"].view(np.uint32), fb.numpy().view(np.uint32))
    assert np.array_equal(got["w_atoms"], wa.numpy()) and np.array_equal(got["w_bonds"], wb.numpy())
    in_ptr, in_idx, out_ptr, out_idx, dst = mpn_oracle.csr_from_a2b(a2b.tolist(), b2a.tolist(), b2revb.tolist())
    nb1 = b2a.numel() - 1
    assert np.array_equal(got["in_ptr"], in_ptr) and np.array_equal(got["in_ptr"], out_ptr)
    assert np.array_equal(got["in_idx"][:nb1], in_idx) and np.array_equal(got["out_idx"][:nb1], out_idx)
    assert np.array_equal(got["src"], b2a.numpy()) and np.array_equal(got["rev"], b2revb.numpy())
    assert np.array_equal(got["dst"], dst)
    assert [(int(s), int(n)) for s, n in zip(a_start, na)] == a_scope
    assert np.array_equal(got["dop"], np.asarray([float(x) for x in dop], np.float32))


def test_store_rejects_graphs_it_cannot_hold():
    bad = synth.make_graphs("copolymer", 2, 5)
    bad[0].f_bonds = bad[0].f_bonds.copy()
    bad[0].f_bonds[1, 3] += 1.0
    with pytest.raises(ValueError):
        DeviceGraphStore.pack(bad)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        DeviceGraphStore(synth.make_graphs("qm9", 2, 0), torch.device("cpu"))
