"""Where does the end-to-end loop at B=4096 spend its wall time: the per-step walls and the final drain, with the packed
upload (one pinned arena) and with the tensor-by-tensor upload."""
import os, sys, time, types
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from polymer_chemprop_b200 import MPN, BatchMolGraph, DeviceGraph, synth  # noqa: E402

DEV = torch.device("cuda", 0)
B = 4096
a = types.SimpleNamespace(hidden=300, depth=3, math="tf32x3")
graphs = synth.make_graphs("copolymer", B, seed=0)
for packed in (True, False, True):
    bmg = BatchMolGraph(graphs)
    if packed:
        bmg.pin_memory()
    else:
        for name in bmg._TENSORS:
            t = getattr(bmg, name)
            if t is not None:
                setattr(bmg, name, t.pin_memory())
    model = MPN(bench.make_model_args(a, DEV), atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(DEV)
    params = [p for p in model.parameters() if p.requires_grad]
    d_out = torch.randn(B, 300, device=DEV)
    host_out = torch.empty(B, 300).pin_memory()
    cs, ds = torch.cuda.Stream(DEV), torch.cuda.Stream(DEV)

    def nxt():
        hg = bmg.fresh_view()
        DeviceGraph.prefetch(hg, DEV, cs)
        return hg

    st = {"n": nxt()}

    def step():
        cur = st["n"]; st["n"] = nxt()
        for p in params: p.grad = None
        out = model([cur])
        ev = torch.cuda.Event(); ev.record()
        with torch.cuda.stream(ds):
            ds.wait_event(ev); host_out.copy_(out.detach(), non_blocking=True)
        out.detach().record_stream(ds)
        out.backward(d_out)
        ds.synchronize()

    for _ in range(int(os.environ.get("WARM", "5"))): step()
    torch.cuda.synchronize()
    t0 = time.perf_counter(); walls = []
    for _ in range(20):
        ts = time.perf_counter(); step(); walls.append(1e3 * (time.perf_counter() - ts))
    t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"packed={packed}: loop {1e3*(t1-t0):.1f} ms, final drain {1e3*(t2-t1):.1f} ms, median step {sorted(walls)[10]:.3f} ms, mem {torch.cuda.memory_reserved()/1e9:.2f} GB reserved", flush=True)
