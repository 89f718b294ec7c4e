"""Where does the host-buffer (e2e) step spend its time?  Wall-clock per phase with synchronisation."""
import os, sys, time, types
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from polymer_chemprop_b200 import MPN, BatchMolGraph, DeviceGraph, synth
import bench

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
dev = torch.device("cuda", 0)
a = types.SimpleNamespace(hidden=300, depth=3, math="tf32")
graphs = synth.make_graphs("copolymer", B, seed=0)
bmg = BatchMolGraph(graphs)
model = MPN(bench.make_model_args(a, dev), atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(dev)
bench.init_weights(model)
comps = bmg.get_components()
pinned = tuple(x.pin_memory() for x in comps[:7]) + tuple(comps[7:])
d_out = torch.randn(B, 300, device=dev)
host_out = torch.empty(B, 300).pin_memory()
sync = torch.cuda.synchronize
for it in range(6):
    t = [time.perf_counter()]
    hg = bench.HostGraph(pinned)
    dg = DeviceGraph.get(hg, dev); sync(); t.append(time.perf_counter())
    out = model([hg]); sync(); t.append(time.perf_counter())
    out.backward(d_out); sync(); t.append(time.perf_counter())
    host_out.copy_(out.detach(), non_blocking=True); sync(); t.append(time.perf_counter())
    for p in model.parameters():
        p.grad = None
    del out, dg, hg
    t.append(time.perf_counter())
    print(f"step {it}: upload+csr {1e3*(t[1]-t[0]):.2f} ms, forward {1e3*(t[2]-t[1]):.2f}, backward {1e3*(t[3]-t[2]):.2f}, "
          f"d2h {1e3*(t[4]-t[3]):.2f}, free {1e3*(t[5]-t[4]):.2f}; reserved {torch.cuda.memory_reserved()/2**30:.2f} GiB")

print("--- prefetch pipeline (bench.py e2e order: step i launched first, then the upload of batch i+1)")
tail = bmg.f_bonds_tail.pin_memory()
copy_stream = torch.cuda.Stream(device=dev)
for use_tail in (False, True):
    def next_batch():
        hg = bench.HostGraph(pinned)
        if use_tail:
            hg.f_bonds_tail = tail
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(copy_stream)
        t0 = time.perf_counter()
        DeviceGraph.prefetch(hg, dev, copy_stream)
        host_ms = 1e3 * (time.perf_counter() - t0)
        e1.record(copy_stream)
        return hg, e0, e1, host_ms
    nxt = next_batch()
    sync()
    for it in range(8):
        t0 = time.perf_counter()
        cur = nxt
        c0 = torch.cuda.Event(enable_timing=True); c1 = torch.cuda.Event(enable_timing=True); c2 = torch.cuda.Event(enable_timing=True)
        c0.record()
        out = model([cur[0]])
        t1 = time.perf_counter()
        out.backward(d_out)
        t2 = time.perf_counter()
        c1.record()
        host_out.copy_(out.detach(), non_blocking=True)
        c2.record()
        nxt = next_batch()
        t3 = time.perf_counter()
        torch.cuda.current_stream().synchronize()
        t4 = time.perf_counter()
        sync()
        for p in model.parameters():
            p.grad = None
        print(f"tail={use_tail} step {it}: wall {1e3*(t4-t0):.2f} ms | host: fwd launch {1e3*(t1-t0):.2f}, bwd launch {1e3*(t2-t1):.2f}, "
              f"prefetch launch {1e3*(t3-t2):.2f}, wait {1e3*(t4-t3):.2f} | gpu: fwd+bwd {c0.elapsed_time(c1):.2f}, d2h {c1.elapsed_time(c2):.2f}, "
              f"copy stream (upload+build of next) {nxt[1].elapsed_time(nxt[2]):.2f}")
