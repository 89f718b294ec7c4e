"""GPU: repeated steps must give the same bits.  The captured backward pass runs its weight-gradient kernels on a
parallel branch of the CUDA graph (csrc/encoder.cu: SideBranch) and every tensor-core kernel is a persistent, warp-
specialised pipeline: a protocol race shows up as ONE step in dozens that differs.  (This is the check that caught the
race of the experimental second epilogue group, profiles/r2_fwd_abl.log section G; scripts/side_branch_stress.py is the
longer form.)  Bitwise equality, no tolerance."""
import pytest
import torch

import _util
from polymer_chemprop_b200 import BatchMolGraph, synth

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)


@pytest.mark.parametrize("mode,depth,undirected,bias", [("bf16x3", 3, False, False), ("bf16x3", 4, True, True),
                                                        ("tf32x3", 3, False, False), ("tf32x3", 4, False, True),
                                                        ("tf32", 5, True, False)])
def test_forty_steps_in_a_row_give_the_same_bits_with_and_without_the_side_branch(mode, depth, undirected, bias, monkeypatch):
    bmg = BatchMolGraph(synth.make_graphs("copolymer", 300, 61, random_edge_weights=True) + [synth.empty_graph()] +
                        synth.make_graphs("qm9", 50, 62))
    torch.manual_seed(1)
    model = _util.MPN(_util.make_args(hidden_size=300, depth=depth, undirected=undirected, bias=bias, wd_math_mode=mode),
                      atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(DEV)
    d_out = torch.randn(len(bmg.a_scope), 300, generator=torch.Generator().manual_seed(2))
    monkeypatch.setenv("WD_SIDE_BRANCH", "0")
    ref_out, ref = _util.run_fwd_bwd(model, bmg, d_out)
    monkeypatch.setenv("WD_SIDE_BRANCH", "1")
    for step in range(40):
        out, grads = _util.run_fwd_bwd(model, bmg, d_out)
        assert torch.equal(out, ref_out), step
        for k in ref:
            assert torch.equal(grads[k], ref[k]), (step, k)
