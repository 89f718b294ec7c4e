"""Where does a batch-50 step spend its time?  (a) the Python step, (b) the two C calls alone on preallocated buffers,
with and without CUDA graphs (WD_GRAPH), (c) cProfile of the Python step."""
import cProfile
import ctypes as C
import os
import pstats
import sys
import time
import types

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from polymer_chemprop_b200 import MPN, BatchMolGraph, DeviceGraph, _capi, synth  # noqa: E402

DEV = torch.device("cuda", 0)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 50
a = types.SimpleNamespace(hidden=300, depth=3, math=os.environ.get("WDMPNN_MATH", "tf32x3"))
bmg = BatchMolGraph(synth.make_graphs("copolymer", B, seed=0))
model = MPN(bench.make_model_args(a, DEV), atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(DEV)
bench.init_weights(model)
params = [p for p in model.parameters() if p.requires_grad]
d_out = torch.randn(B, 300, device=DEV)


def step():
    for p in params:
        p.grad = None
    out = model([bmg])
    out.backward(d_out)


def wall(fn, n=300):
    for _ in range(30):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    t_host = time.perf_counter() - t0
    torch.cuda.synchronize()
    return 1e6 * t_host / n, 1e6 * (time.perf_counter() - t0) / n


for g in ("1", "0"):
    os.environ["WD_GRAPH"] = g
    h, t = wall(step)
    print(f"WD_GRAPH={g}: python step  host {h:7.1f} us  total {t:7.1f} us per step")

# (b) the C calls alone
enc = model.encoder[0]
dg = DeviceGraph.get(bmg, DEV)
L, p = _capi.lib(), _capi.ptr
for g in (0, 2):
    cfg = enc._config(dg)
    cfg.flags = g
    nbytes = L.wd_workspace_bytes(C.byref(cfg), dg.n_atoms, dg.n_bonds, dg.n_mols)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=DEV)
    out = torch.empty(B, 300, device=DEV)
    W = [enc.W_i.weight, None, enc.W_h.weight, None, enc.W_o.weight, enc.W_o.bias]
    weights = _capi.WdWeights(*[p(t) for t in W])
    G = [torch.empty_like(t) if t is not None else None for t in W]
    grads = _capi.WdGrads(*[p(t) for t in G])
    s = _capi.current_stream(DEV)

    def ccalls():
        L.wd_encoder_forward(C.byref(dg.struct), C.byref(weights), C.byref(cfg), p(ws), nbytes, p(out), s)
        L.wd_encoder_backward(C.byref(dg.struct), C.byref(weights), C.byref(cfg), p(ws), nbytes, p(d_out), C.byref(grads), s)

    h, t = wall(ccalls)
    print(f"flags={g} ({'no graph' if g else 'graph'}): C calls only  host {h:7.1f} us  total {t:7.1f} us per step")

    def fwd_only():
        L.wd_encoder_forward(C.byref(dg.struct), C.byref(weights), C.byref(cfg), p(ws), nbytes, p(out), s)
    h, t = wall(fwd_only)
    print(f"           forward only  host {h:7.1f} us  total {t:7.1f} us")

os.environ["WD_GRAPH"] = "1"
pr = cProfile.Profile()
pr.enable()
for _ in range(300):
    step()
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
