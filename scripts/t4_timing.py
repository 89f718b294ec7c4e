"""Per-role wait-cycle breakdown of the gathered depth-step kernel (k_tc_gemm4) inside a real encoder forward.
Needs a library built with -DWD_T2_TIMING -DWD_T2_TIMING_ONLY=4 (WDMPNN_LIB=.../libwdmpnn_timing.so)."""
import ctypes as C
import os
import sys
import types

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from polymer_chemprop_b200 import MPN, BatchMolGraph, _capi, synth  # noqa: E402

NAMES = {0: "xform <- raw_full (loads landed)", 2: "loader <- a_empty (stage free)", 3: "bload <- b_empty", 4: "mma <- tmem_empty",
         5: "mma <- op_full (transform done)", 6: "mma <- b_full", 7: "epi <- tmem_full", 8: "epi <- r_full", 9: "rload <- r_empty"}
dev = torch.device("cuda", 0)
a = types.SimpleNamespace(hidden=300, depth=3, math="tf32")
bmg = BatchMolGraph(synth.make_graphs("copolymer", int(sys.argv[1]) if len(sys.argv) > 1 else 4096, seed=0))
model = MPN(bench.make_model_args(a, dev), atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(dev)
bench.init_weights(model)
L = _capi.lib()
buf = (C.c_uint64 * 16)()
with torch.no_grad():
    for it in range(3):
        L.wd_debug_timing(buf, 1)
        model([bmg])
        torch.cuda.synchronize()
L.wd_debug_timing(buf, 0)
tot = buf[15] or 1
print(f"WD_T4_X1_TMA={os.environ.get('WD_T4_X1_TMA', '1')}: CTA 0, {model.encoder[0].depth - 1} launches of k_tc_gemm4, {tot} cycles in total")
for i, n in NAMES.items():
    print(f"   {n:36s} {100.0 * buf[i] / tot:6.1f} % of the kernel's cycles waiting")
