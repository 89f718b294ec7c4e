"""CPU: the drop-in BatchMolGraph reproduces the reference BatchMolGraph's tensors bit-exactly
(goldens were built by the reference's own BatchMolGraph.__init__, featurization.py:757-813)."""
import numpy as np
import pytest
import torch

import _golden
from oracle import make_golden, ref_shim
from polymer_chemprop_b200 import BatchMolGraph


@pytest.mark.parametrize("name", _golden.NAMES)
def test_batch_graph_bit_exact_vs_golden(name):
    build, _cfg = make_golden._cases()[name]
    bmg = BatchMolGraph(build())
    ref = _golden.load(name)["components"]
    got = bmg.get_components()
    for i in range(7):
        assert got[i].dtype == ref[i].dtype and got[i].shape == ref[i].shape, i
        assert torch.equal(got[i], ref[i]), i
    assert got[7] == ref[7] and got[8] == ref[8]
    assert [float(x) for x in got[9]] == ref[9]


@pytest.mark.skipif(not ref_shim.reference_available(), reason="reference tree absent")
def test_batch_graph_vs_live_reference():
    from polymer_chemprop_b200 import synth
    graphs = synth.make_graphs("copolymer", 5, 123, random_edge_weights=True) + [synth.empty_graph()]
    ref = ref_shim.load_reference().BatchMolGraph([g.as_reference_lists() for g in graphs])
    mine = BatchMolGraph(graphs)
    for a, b in zip(mine.get_components()[:7], ref.get_components()[:7]):
        assert torch.equal(a, b)
    assert mine.a_scope == ref.a_scope and mine.b_scope == ref.b_scope
    assert mine.max_num_bonds == ref.max_num_bonds and mine.n_atoms == ref.n_atoms
    assert torch.equal(mine.get_a2a(), ref.get_a2a()) and torch.equal(mine.get_b2b(), ref.get_b2b())


def test_synthetic_shapes():
    """copolymer-shaped ~23 atoms / ~66 directed bonds, K 7-8, self-loop pairs."""
    from polymer_chemprop_b200 import synth
    bmg = BatchMolGraph(synth.make_graphs("copolymer", 200, 0))
    assert 20 < (bmg.n_atoms - 1) / 200 < 26 and 58 < (bmg.n_bonds - 1) / 200 < 72
    assert bmg.max_num_bonds in (7, 8)
    dst = bmg.b2a[bmg.b2revb]
    assert int((dst[1:] == bmg.b2a[1:]).sum()) == 200 * 8       # 4 self pairs x 2 directed bonds
    assert torch.equal(bmg.f_bonds[:, :133], bmg.f_atoms[bmg.b2a])
    q = BatchMolGraph(synth.make_graphs("qm9", 200, 0))
    assert q.max_num_bonds <= 4 and 7 <= (q.n_atoms - 1) / 200 <= 9


def test_bond_tail_only_when_prefix_property_holds():
    """BatchMolGraph.f_bonds_tail (the compact upload form, csrc/features.cu) exists iff every molecule's bond
    rows start with the source atom's row bit for bit (featurization.py:467-468 builds them that way)."""
    from polymer_chemprop_b200 import synth
    graphs = synth.make_graphs("copolymer", 20, 7, dense_features=True) + [synth.empty_graph()] + \
        synth.make_graphs("qm9", 5, 8)
    bmg = BatchMolGraph(graphs)
    assert bmg.f_bonds_tail is not None and bmg.f_bonds_tail.is_contiguous()
    assert torch.equal(bmg.f_bonds_tail, bmg.f_bonds[:, bmg.atom_fdim:])
    assert torch.equal(torch.cat([bmg.f_atoms[bmg.b2a], bmg.f_bonds_tail], dim=1), bmg.f_bonds)
    bad = synth.make_graphs("copolymer", 3, 9)
    bad[1].f_bonds = bad[1].f_bonds.copy()
    bad[1].f_bonds[2, 5] += 1.0                                   # one bond row no longer repeats its atom row
    assert BatchMolGraph(bad).f_bonds_tail is None
    neg0 = synth.make_graphs("copolymer", 2, 10)
    neg0[0].f_bonds = neg0[0].f_bonds.copy()
    neg0[0].f_atoms = neg0[0].f_atoms.copy()
    a = int(neg0[0].b2a[0])
    col = int(np.argmin(neg0[0].f_atoms[a]))                      # a 0.0 entry: -0.0 == 0.0 numerically, not bitwise
    assert neg0[0].f_atoms[a, col] == 0.0
    neg0[0].f_bonds[0, col] = -0.0
    assert BatchMolGraph(neg0).f_bonds_tail is None


def test_pack_arena_keeps_values_dtypes_and_shapes():
    """pack_arena lays the upload set out in one host buffer (256-byte aligned offsets) and re-points the attributes
    at views of it; nothing else changes (CPU here: the arena is not pinned)."""
    import torch
    from polymer_chemprop_b200 import BatchMolGraph, synth
    graphs = synth.make_graphs("copolymer", 9, 3, dense_features=True)
    a, b = BatchMolGraph(graphs), BatchMolGraph(graphs).pack_arena(pin=False)
    assert b._arena.dtype == torch.uint8 and set(b._arena_layout) == set(BatchMolGraph._UPLOAD)
    end = 0
    for name in BatchMolGraph._UPLOAD:
        x, y = getattr(a, name), getattr(b, name)
        off, nb = b._arena_layout[name]
        assert off % 256 == 0 and off >= end and nb == x.numel() * x.element_size()
        end = off + nb
        assert x.dtype == y.dtype and x.shape == y.shape and torch.equal(x, y), name
        assert y.data_ptr() == b._arena.data_ptr() + off
    assert end <= b._arena.numel()
    assert [torch.equal(u, v) for u, v in zip(a.get_components()[:7], b.get_components()[:7])] == [True] * 7
    assert b.pack_arena(pin=False) is b and b.fresh_view()._arena is b._arena          # idempotent, shared by views


def test_bucketed_capacities_round_up_on_a_coarse_grid_and_leave_small_sizes_alone():
    """_capi.bucket / empty_roomy: what the workspace and the store-assembled batches are allocated with, so that batches
    whose sizes differ by a fraction of a per cent reuse the same cached blocks."""
    import torch
    from polymer_chemprop_b200 import _capi
    assert [_capi.bucket(n) for n in (0, 1, 4096, (1 << 20) - 1)] == [0, 1, 4096, (1 << 20) - 1]
    for n in (1 << 20, (1 << 20) + 1, 3_778_720_256, 81_000_000, 81_300_000):
        b = _capi.bucket(n)
        assert n <= b <= n * 1.0626 and b % (1 << (n.bit_length() - 5)) == 0
    assert _capi.bucket(81_000_000) == _capi.bucket(81_300_000)              # 0.4 % apart: one bucket
    t = _capi.empty_roomy((300_000, 4), torch.float32, "cpu")
    assert t.shape == (300_000, 4) and t.is_contiguous() and t.untyped_storage().nbytes() == 4 * _capi.bucket(1_200_000)
    assert _capi.empty_roomy((7,), torch.int32, "cpu").untyped_storage().nbytes() == 28


def test_high_water_capacities_grow_monotonically_and_ignore_much_smaller_requests():
    import torch
    from polymer_chemprop_b200 import _capi
    H, key = _capi.empty_high_water, ("test", 1)
    nbytes = lambda t: t.untyped_storage().nbytes()
    a = H(key, (81_000_000,), torch.uint8, "cpu")          # first request: its bucket with 3 % headroom
    b = H(key, (80_000_000,), torch.uint8, "cpu")          # smaller, same order of magnitude: the capacity so far
    c = H(key, (82_000_000,), torch.uint8, "cpu")          # 1.2 % larger: still inside the headroom
    d = H(key, (90_000_000,), torch.uint8, "cpu")          # a new maximum beyond it: the capacity grows
    e = H(key, (81_000_000,), torch.uint8, "cpu")
    f = H(key, (1_000,), torch.uint8, "cpu")               # another workload: exact
    cap0 = _capi.bucket(81_000_000 + 81_000_000 // 32)
    assert nbytes(a) == nbytes(b) == nbytes(c) == cap0 and 82_000_000 <= cap0 <= 81_000_000 * 1.1
    assert nbytes(d) == nbytes(e) == _capi.bucket(90_000_000 + 90_000_000 // 32) > cap0
    assert nbytes(f) == 1_000
    assert [t.numel() for t in (a, b, c, d, e, f)] == [81_000_000, 80_000_000, 82_000_000, 90_000_000, 81_000_000, 1_000]
