"""Stress of the captured backward pass with its side branch (csrc/encoder.cu: SideBranch): 60 steps per depth, every
gradient compared bitwise with the step run without the side branch (WD_SIDE_BRANCH=0)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import _util  # noqa: E402
from polymer_chemprop_b200 import BatchMolGraph, synth  # noqa: E402

DEV = torch.device("cuda", 0)
bmg = BatchMolGraph(synth.make_graphs("copolymer", 300, 61, random_edge_weights=True) + [synth.empty_graph()] +
                    synth.make_graphs("qm9", 50, 62))
total_bad = 0
for mode, depth, und, bias in (("bf16x3", 3, False, False), ("bf16x3", 5, True, True), ("tf32x3", 3, False, False), ("tf32x3", 4, False, True), ("tf32", 5, True, False), ("fp32", 3, True, True)):
    torch.manual_seed(1)
    model = _util.MPN(_util.make_args(hidden_size=300, depth=depth, undirected=und, bias=bias, wd_math_mode=mode),
                      atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(DEV)
    d_out = torch.randn(len(bmg.a_scope), 300, generator=torch.Generator().manual_seed(2))
    os.environ["WD_SIDE_BRANCH"] = "0"
    ref_out, ref = _util.run_fwd_bwd(model, bmg, d_out)
    os.environ["WD_SIDE_BRANCH"] = os.environ.get("STRESS_SB", "1")
    bad = {}
    for it in range(60):
        out, g = _util.run_fwd_bwd(model, bmg, d_out)
        for k in g:
            if not torch.equal(g[k], ref[k]):
                bad[k] = bad.get(k, 0) + 1
    total_bad += sum(bad.values())
    print(mode, "depth", depth, "undirected", und, "bias", bias, "-> mismatching steps per tensor out of 60:", bad or "none", flush=True)
print("SIDE BRANCH STRESS:", "FAIL" if total_bad else "OK")
sys.exit(1 if total_bad else 0)
