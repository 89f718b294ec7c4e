"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz from the UNMODIFIED reference.

Run in the build container (needs /root/reference):  python -m oracle.make_golden
Each fixture stores: the reference BatchMolGraph tensors built by the reference's own
BatchMolGraph.__init__ from duck-typed MolGraphs, the state_dict, d_out, the reference
MPN output and the autograd gradients, plus the config.  The GPU box has no reference
tree; tests there read only these files.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle import ref_shim  # noqa: E402
from polymer_chemprop_b200 import synth  # noqa: E402

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")

# name -> (graph builder, encoder config)
def _cases():
    def g(kind, n, seed, **kw):
        return lambda: synth.make_graphs(kind, n, seed, dense_features=True, **kw)

    def edge():
        rng = np.random.default_rng(7)
        gs = synth.make_graphs("copolymer", 2, 11, dense_features=True, random_edge_weights=True)
        return [gs[0], synth.empty_graph(), synth.single_atom_graph(rng), gs[1], synth.qm9_graph(rng, True)]

    def singles():
        rng = np.random.default_rng(3)
        return [synth.single_atom_graph(rng) for _ in range(3)]

    full = dict(hidden_size=300, depth=3)
    base = dict(hidden_size=100, depth=3)   # small h keeps the committed fixtures small
    return {
        "copoly_b2": (g("copolymer", 2, 1), dict(full)),
        "copoly_b50": (lambda: synth.make_graphs("copolymer", 50, 0), dict(full)),
        "copoly_b8_randw": (g("copolymer", 8, 2, random_edge_weights=True), dict(base)),
        "qm9_b16": (g("qm9", 16, 3), dict(base)),
        "copoly_b4_sum": (g("copolymer", 4, 4), dict(base, aggregation="sum")),
        "copoly_b4_norm": (g("copolymer", 4, 5), dict(base, aggregation="norm", aggregation_norm=37)),
        "copoly_b4_undirected": (g("copolymer", 4, 6, random_edge_weights=True), dict(base, undirected=True)),
        "copoly_b4_bias": (g("copolymer", 4, 7), dict(base, bias=True)),
        "copoly_b3_d5_h64": (g("copolymer", 3, 8), dict(hidden_size=64, depth=5)),
        "copoly_b3_d2": (g("copolymer", 3, 9), dict(hidden_size=52, depth=2)),
        "edge_cases": (edge, dict(base)),
        "all_single_atoms": (singles, dict(base)),
    }


def run_reference(graphs, cfg, seed=0):
    ref = ref_shim.load_reference()
    bmg = ref.BatchMolGraph([g.as_reference_lists() for g in graphs])
    args = ref_shim.make_args(**cfg)
    torch.manual_seed(seed)
    model = ref.MPN(args, atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim)
    for name, p in model.named_parameters():          # nn_utils.py:102-112 semantics + non-zero biases
        if not p.requires_grad:
            continue
        if p.dim() == 1:
            torch.nn.init.normal_(p, 0.0, 0.1)
        else:
            torch.nn.init.xavier_normal_(p)
    out = model([bmg])
    d_out = torch.randn(out.shape, generator=torch.Generator().manual_seed(seed + 1))
    out.backward(d_out)
    enc = model.encoder[0]
    sd = {k: v.detach().numpy() for k, v in enc.state_dict().items()}
    grads = {k: p.grad.numpy() for k, p in enc.named_parameters() if p.grad is not None}
    return bmg, sd, d_out.numpy(), out.detach().numpy(), grads


def main():
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    for name, (build, cfg) in _cases().items():
        graphs = build()
        bmg, sd, d_out, out, grads = run_reference(graphs, cfg)
        fa, fb, wa, wb, a2b, b2a, b2revb, a_scope, b_scope, dop = bmg.get_components()
        blob = dict(f_atoms=fa.numpy(), f_bonds=fb.numpy(), w_atoms=wa.numpy(), w_bonds=wb.numpy(),
                    a2b=a2b.numpy(), b2a=b2a.numpy(), b2revb=b2revb.numpy(),
                    a_scope=np.asarray(a_scope, np.int64).reshape(-1, 2),
                    b_scope=np.asarray(b_scope, np.int64).reshape(-1, 2),
                    degree_of_polym=np.asarray([float(x) for x in dop], np.float64),
                    d_out=d_out, out=out,
                    cfg_keys=np.asarray(list(cfg.keys())), cfg_vals=np.asarray([str(v) for v in cfg.values()]))
        if name == "copoly_b50":                       # sparse features compress well; keep fixture small
            pass
        for k, v in sd.items():
            blob["param:" + k] = v
        for k, v in grads.items():
            blob["grad:" + k] = v
        path = os.path.join(GOLDEN_DIR, name + ".npz")
        np.savez_compressed(path, **blob)
        print(f"{name}: N_a={fa.shape[0]} N_b={fb.shape[0]} K={a2b.shape[1]} G={len(a_scope)} "
              f"-> {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
