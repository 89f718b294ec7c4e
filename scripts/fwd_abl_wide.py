"""Forward-only timing of the wide configuration (h = 2400, d = 6) for one library build (WDMPNN_LIB selects it; ablation
builds carry -DWD_ABL=<bits> and give wrong results by design).  Prints the encoder forward time and the cost of one
depth step (difference of the depth-d and the depth-2 forward)."""
import os
import sys
import types

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from polymer_chemprop_b200 import MPN, BatchMolGraph, synth  # noqa: E402

DEV = torch.device("cuda", 0)


def timed(fn, n=4):
    for _ in range(2):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


if __name__ == "__main__":
    hidden, depth, batch = (int(x) for x in (sys.argv[1:4] + ["2400", "6", "2048"][len(sys.argv) - 1:]))
    tag = os.path.basename(os.environ.get("WDMPNN_LIB", "libwdmpnn.so"))
    bmg = BatchMolGraph(synth.make_graphs("copolymer", batch, seed=0))
    out = [f"{tag} h={hidden} d={depth} B={batch}:"]
    for mode in ("tf32x3", "tf32", "bf16x3"):
        t = {}
        for d in (depth, 2):
            a = types.SimpleNamespace(hidden=hidden, depth=d, math=mode)
            model = MPN(bench.make_model_args(a, DEV), atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(DEV)
            bench.init_weights(model)

            def fwd():
                with torch.no_grad():
                    model([bmg])
            t[d] = timed(fwd)
        out.append(f"{mode} fwd {t[depth]:.2f} ms, depth step {(t[depth] - t[2]) / (depth - 2):.3f} ms")
    print(" | ".join(out), flush=True)
