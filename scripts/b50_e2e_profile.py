"""Where does an END-TO-END batch-50 step (pinned host batch -> result on the host) spend its host time?  The loop of
bench.py's `e2e` leg, timed piece by piece and under cProfile."""
import cProfile
import os
import pstats
import sys
import time
import types

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from polymer_chemprop_b200 import MPN, BatchMolGraph, DeviceGraph, synth  # noqa: E402

DEV = torch.device("cuda", 0)
B = int(sys.argv[1]) if len(sys.argv) > 1 else 50
a = types.SimpleNamespace(hidden=300, depth=3, math=os.environ.get("WDMPNN_MATH", "tf32x3"))
bmg = BatchMolGraph(synth.make_graphs("copolymer", B, seed=0)).pin_memory()
model = MPN(bench.make_model_args(a, DEV), atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(DEV)
bench.init_weights(model)
params = [p for p in model.parameters() if p.requires_grad]
d_out = torch.randn(B, 300, device=DEV)
host_out = torch.empty(B, 300).pin_memory()
copy_stream, d2h_stream = torch.cuda.Stream(DEV), torch.cuda.Stream(DEV)
T = {"prefetch": 0.0, "forward": 0.0, "d2h": 0.0, "backward": 0.0, "wait": 0.0}


def next_batch():
    hg = bmg.fresh_view()
    DeviceGraph.prefetch(hg, DEV, copy_stream)
    return hg


state = {"next": next_batch()}


def e2e_step():
    t0 = time.perf_counter()
    cur = state["next"]
    state["next"] = next_batch()
    t1 = time.perf_counter()
    for p in params:
        p.grad = None
    out = model([cur])
    t2 = time.perf_counter()
    fwd_done = torch.cuda.Event()
    fwd_done.record()
    with torch.cuda.stream(d2h_stream):
        d2h_stream.wait_event(fwd_done)
        host_out.copy_(out.detach(), non_blocking=True)
    out.detach().record_stream(d2h_stream)
    t3 = time.perf_counter()
    out.backward(d_out)
    t4 = time.perf_counter()
    d2h_stream.synchronize()
    t5 = time.perf_counter()
    for k, v in zip(T, (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4)):
        T[k] += v


for _ in range(30):
    e2e_step()
torch.cuda.synchronize()
for k in T:
    T[k] = 0.0
n = 300
t0 = time.perf_counter()
for _ in range(n):
    e2e_step()
torch.cuda.synchronize()
tot = time.perf_counter() - t0
print(f"B={B}: {1e6 * tot / n:.1f} us per step = {B * n / tot:.0f} graphs/s; host time per piece (us): "
      + ", ".join(f"{k} {1e6 * v / n:.1f}" for k, v in T.items()))
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    e2e_step()
pr.disable()
torch.cuda.synchronize()
pstats.Stats(pr).sort_stats("cumulative").print_stats(45)
