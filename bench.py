"""Benchmark of the wD-MPNN encoder hot path (BASELINE.json metric: encoder graphs/sec, fwd+bwd,
h=300, d=3, synthetic copolymer-shaped graphs, batch 4096 per GPU).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A step = MPN.forward(batch) + out.backward(d_out) on one batch (+ one NCCL all-reduce of the
flat gradient when N > 1).  Rank 0 prints ONE JSON line.  See DESIGN.md "Measurement".
"""
import argparse
import ctypes as C
import gc
import json
import os
import subprocess
import sys
import threading
import time
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "wD-MPNN encoder graphs/sec (fwd+bwd) at h=300,d=3"
UNIT = "graphs/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=4096, help="graphs per GPU per step")
    ap.add_argument("--hidden", type=int, default=300)
    ap.add_argument("--depth", type=int, default=3)
    ap.add_argument("--kind", default="copolymer", choices=["copolymer", "qm9"])
    ap.add_argument("--math", default=os.environ.get("WDMPNN_MATH", "bf16x3"), choices=["fp32", "tf32", "tf32x3", "bf16x3"],
                    help="bf16x3 (default, the drop-in's default): tcgen05 tensor cores, error-compensated forward "
                         "contractions split at bfloat16 (encoder output <= 2e-5, predictions <= 1e-4); tf32x3: the same "
                         "split at TF32 (<= 1e-5); tf32: single pass (<= 1e-3 on the encoder output); fp32: CUDA-core "
                         "contractions")
    ap.add_argument("--cpu-sample", type=int, default=512, help="graphs in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-small-batch", action="store_true", help="skip the batch-50 sub-run (key small_batch)")
    ap.add_argument("--no-single-pass", action="store_true", help="skip the single-pass TF32 comparison (key single_pass_tf32)")
    ap.add_argument("--store-graphs", type=int, default=None,
                    help="graphs per rank in the device-resident store (default: the batch at N=1; 125000 at N>1 = "
                         "BASELINE configs[4]'s 10^6 graphs over 8 GPUs; distinct synthetic graphs are repeated)")
    return ap.parse_args()


def make_model_args(a, device):
    return types.SimpleNamespace(
        hidden_size=a.hidden, depth=a.depth, bias=False, dropout=0.0, activation="ReLU",
        atom_messages=False, undirected=False, aggregation="mean", aggregation_norm=100,
        atom_descriptors=None, atom_descriptors_size=0, device=device, mpn_shared=False,
        number_of_molecules=1, features_only=False, use_input_features=False,
        overwrite_default_atom_features=False, overwrite_default_bond_features=False,
        wd_math_mode=a.math)


def init_weights(model, seed=0):
    """initialize_weights semantics (chemprop/nn_utils.py:102-112): xavier-normal matrices, zero vectors."""
    torch.manual_seed(seed)
    for p in model.parameters():
        if not p.requires_grad:
            continue
        if p.dim() == 1:
            torch.nn.init.constant_(p, 0)
        else:
            torch.nn.init.xavier_normal_(p)


def workload_name(a):
    if a.kind == "qm9":
        cfg = "configs[2]: unweighted D-MPNN on QM9-shaped graphs"
    elif a.hidden == 2400 and a.depth == 6:
        cfg = "configs[3]: wide/deep, GEMM-dominated"
    elif a.hidden == 300 and a.depth == 3:
        cfg = "configs[1]" if world_size() == 1 else "configs[4]: data parallel, 4096 graphs per GPU"
    else:
        cfg = "not a BASELINE.json configuration"
    return (f"synthetic {a.kind}-shaped graphs, batch {a.batch}/GPU, hidden {a.hidden}, depth {a.depth} "
            f"(BASELINE.json {cfg})")


def world_size():
    return int(os.environ.get("WORLD_SIZE", "1"))


# --------------------------------------------------------------------------------------------
# CPU baseline: the oracle port of the reference's op sequence on the host cores
# --------------------------------------------------------------------------------------------
def cpu_port_time(graphs, a, n, threads, steps, warmup, readout=True, budget_s=60.0):
    """Times `steps` forward + backward passes (after `warmup`) of the reference's op sequence (oracle/mpn_oracle.py,
    torch CPU fp32) on the first n graphs with `threads` torch threads.  Stops early once budget_s is spent (the
    steps actually timed are reported)."""
    from oracle import mpn_oracle
    from polymer_chemprop_b200 import BatchMolGraph
    n = min(n, len(graphs))
    bmg = BatchMolGraph(graphs[:n])
    torch.set_num_threads(threads)
    h, fa, fb = a.hidden, bmg.atom_fdim, bmg.bond_fdim
    gen = torch.Generator().manual_seed(0)
    weights = {"W_i.weight": torch.randn(h, fb, generator=gen) * 0.07,
               "W_h.weight": torch.randn(h, h, generator=gen) * 0.06,
               "W_o.weight": torch.randn(h, fa + h, generator=gen) * 0.05,
               "W_o.bias": torch.zeros(h)}
    d_out = torch.randn(n, h, generator=gen)
    comps = bmg.get_components()
    times, t_begin = [], time.perf_counter()
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        mpn_oracle.encoder_forward_backward(comps, weights, d_out, depth=a.depth, readout=readout)
        dt = time.perf_counter() - t0
        if i >= warmup:
            times.append(dt)
        if time.perf_counter() - t_begin > budget_s and times:
            break
    ms = 1e3 * float(np.mean(times))
    return {"value": n / (ms * 1e-3), "unit": UNIT, "graphs": n, "threads": torch.get_num_threads(), "ms_per_step": ms,
            "steps": len(times), "warmup": warmup, "readout": readout}


def cpu_baseline(graphs, a, steps=10, warmup=2, extras=True, full_batch=False):
    """The reference's CPU implementation of the path on the box's host cores (the oracle PORT: the reference is pure
    Python and /root/reference is not on the GPU box).  Headline = all cores on a CPU_SAMPLE-graph sample of the same
    workload, the most CPU-favourable batch size (the reference's per-molecule readout backward is super-linear in the
    batch: 43 graphs/s at B=4096 in the survey container).  extras: the other figures BASELINE.md section 3 promises --
    batch 50, one thread, message passing only (readout excluded) -- and, for --impl reference, one step at the full
    labelled batch."""
    cores = os.cpu_count() or 1
    n = min(a.cpu_sample, len(graphs))
    head = cpu_port_time(graphs, a, n, cores, steps, warmup, budget_s=120.0)
    out = {"value": head["value"], "unit": UNIT, "cores": cores, "kind": "port", "threads": head["threads"],
           "ms_per_step": head["ms_per_step"], "steps": head["steps"], "warmup": head["warmup"],
           "sample": f"first {n} graphs of the rank-0 batch per step (forward + backward), reference op sequence "
                     f"(oracle/mpn_oracle.py) on torch CPU fp32, {cores} threads; the cost per graph grows with the batch "
                     f"(per-molecule readout backward), so this sample size FAVOURS the CPU against the full batch"}
    if extras:
        ex = {}
        ex["batch_50_all_cores"] = cpu_port_time(graphs, a, 50, cores, 10, 2, budget_s=10.0)
        ex["one_thread"] = cpu_port_time(graphs, a, n, 1, 2, 1, budget_s=30.0)
        ex["message_passing_only_all_cores"] = cpu_port_time(graphs, a, n, cores, 5, 1, readout=False, budget_s=15.0)
        if full_batch and len(graphs) > n:
            ex["full_batch_all_cores"] = cpu_port_time(graphs, a, len(graphs), cores, 1, 0, budget_s=1.0)
        out["other_figures"] = ex
    torch.set_num_threads(cores)
    return out


def base_config(a, world, G, Na, Nb):
    """The workload description shared by both arms (the driver compares the arms' `config`)."""
    return {"workload": workload_name(a), "global_batch": world * G, "n_atoms": Na, "n_bonds": Nb,
            "parallelism": f"dp{world}" if world > 1 else "single GPU",
            "l2": "per-step working set (workspace >= 2 GB at batch 4096) exceeds the 126 MB L2"}


def run_reference(a):
    """--impl reference: the reference's own CPU implementation of the path, on the arm's config / metric / unit.
    /root/reference is not on the GPU box and it is pure Python, so the timed code is the oracle port of its op
    sequence; each of the K steps (after W warm-ups, as asked) is a bounded sample of the workload -- see
    cpu_baseline()."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from polymer_chemprop_b200 import BatchMolGraph, synth
    graphs = synth.make_graphs(a.kind, a.batch, seed=0)
    bmg = BatchMolGraph(graphs)
    warmup = max(a.warmup, 3)
    cb = cpu_baseline(graphs, a, steps=max(1, a.steps), warmup=warmup, extras=True,
                      full_batch=os.environ.get("WD_REF_FULL", "1") != "0")
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": a.gpus,
            "steps": cb["steps"], "warmup": warmup, "ms_per_step": cb["ms_per_step"],
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": base_config(a, max(1, a.gpus), len(graphs), bmg.n_atoms, bmg.n_bonds),
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------
# clocks during the timed region (B200_PROFILING.md "clocks line")
# --------------------------------------------------------------------------------------------
class ClockSampler:
    """nvidia-smi polled every 20 ms from before the warm-up; only the lines that arrive between begin() and end()
    -- the timed passes -- count (the process takes ~0.1 s to start, longer than a short timed region)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines, self.t0, self.t1 = index, None, [], None, None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.perf_counter(), line.strip()))

    def begin(self):
        self.t0 = time.perf_counter()

    def end(self):
        self.t1 = time.perf_counter()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.proc.terminate()
        self.thread.join(timeout=2)
        inside = [ln for t, ln in self.lines if self.t0 is not None and self.t0 <= t <= (self.t1 or t)]
        window = "timed passes"
        if not inside:                      # region shorter than one poll: the nearest samples taken under load
            inside, window = [ln for _t, ln in self.lines[-5:]], "last samples before the end of the timed passes"
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in inside:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm), "window": window}


# --------------------------------------------------------------------------------------------
# algorithmic bytes per launch of each kernel category (DESIGN.md "Kernels and rooflines")
# --------------------------------------------------------------------------------------------
def algorithmic_bytes(Na, Nb, G, h, Fa, Fb, d, sign_bits=False):
    """Compulsory HBM bytes per launch of each kernel category (average over the launches of the category in one
    step), for the algorithm as implemented: S_t is written by the forward depth steps and read back by dW_h.
    sign_bits: the forward epilogues leave one bit per element of I / M_t (TF32 mode, ReLU) and the message gradient
    reads those (B words per launch) instead of the activation (P)."""
    P, Q = 4.0 * Nb * h, 4.0 * Na * h            # one [N_b,h] / [N_a,h] fp32 pass
    B = 4.0 * Nb * ((h + 31) // 32) if sign_bits else 0.0
    A = B if sign_bits else P                    # what the message gradient reads for act'
    steps = max(d - 1, 1)
    depth_total = ((3 * P + Q + B) + (d - 2) * (4 * P + Q + B)) if d > 1 else 0.0    # t=1 reads I once; every step writes S_t
    if d == 3:     # bench.py's models have no W_h bias: the middle step leaves the running sum to the last one (encoder.cu: defer_sum)
        msg_total = (P + A + Q) + (2 * P + A + Q) + (4 * P + A + Q)
    elif d > 1:
        msg_total = (P + A + Q) + (d - 2) * (4 * P + A + Q) + (3 * P + A + Q)     # first / middle / last step
    else:
        msg_total = P + A + Q
    dw_total = (2 * Q + 4.0 * Na * Fa) + (d - 1) * 2 * P + (P + 4.0 * Nb * Fb)   # dW_o, dW_h x (d-1), dW_i
    dx_total = 2 * Q + (d - 1) * 2 * P                                          # dA_d, dS x (d-1)
    return {
        "input_proj": 4.0 * Nb * Fb + P + B + 4.0 * h * Fb,
        "aggregate": P + Q + 12.0 * Nb + 4.0 * Na,
        "depth_step": depth_total / steps + 8.0 * Nb + 4.0 * h * h,
        "atom_update": 4.0 * Na * Fa + 2 * Q + 4.0 * h * (Fa + h),
        "readout": (2 * Q + 4.0 * G * h) * 0.5 + (Q + 4.0 * G * h) * 0.5,   # bwd (H, dZo, d_out) / fwd (H, out)
        "msg_grad": msg_total / d + 12.0 * Nb,
        "dw": dw_total / (d + 1),
        "dx": dx_total / d + 4.0 * h * h,
        "scatter": P + Q + 8.0 * Nb,
        "misc": 0.0,
        "reduce": 0.0,
    }


def contract_bytes(Na, Nb, G, h, Fa, Fb, d):
    """SURVEY.md section 8(d) "Algorithmic bytes": compulsory traffic as the scope contract counts it (fp32 storage),
    independent of how this implementation splits the work into kernels.  P = one [N_b, h] pass.
    Depth step (fused aggregate + gather + W_h + residual + ReLU): 3P (2P for t = 1) + 16 N_b + 4 N_a + 4 h^2;
    whole step: fwd (3d-2) P + bwd (6d-3) P + 8 N_b F_b + O(N_a h)."""
    P, Q = 4.0 * Nb * h, 4.0 * Na * h
    steps = max(d - 1, 1)
    depth_avg = ((2 * P) + (d - 2) * 3 * P) / steps + 16.0 * Nb + 4.0 * Na + 4.0 * h * h if d > 1 else 0.0
    step = ((3 * d - 2) + (6 * d - 3)) * P + 8.0 * Nb * Fb + 8.0 * Na * Fa + 6 * Q + 8.0 * G * h
    return {"depth_step": depth_avg, "step": step}


def contract_flops(Na, Nb, h, Fa, Fb, d):
    """SURVEY.md section 8(d) "Algorithmic flops" (padding and the 3xTF32 correction passes are NOT credited)."""
    fwd = 2.0 * h * (Nb * Fb + (d - 1) * Nb * h + Na * (Fa + h))
    bwd = 2.0 * h * (Nb * Fb + 2 * (d - 1) * Nb * h + Na * (Fa + h) + Na * h)
    return {"step": fwd + bwd, "depth_step": 2.0 * Nb * h * h}


def measure_tf32_peak(dev):
    """Dense TF32 tensor throughput the way MEASURED_PEAKS.json measures bf16: torch.matmul (cuBLAS) 8192^3 with
    allow_tf32, best of 10, CUDA events.  A yardstick for the tensor-bound configuration only; not on the product path."""
    prev = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        n = 8192
        x, y = torch.randn(n, n, device=dev), torch.randn(n, n, device=dev)
        best = 1e9
        for i in range(12):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch.matmul(x, y)
            e1.record()
            torch.cuda.synchronize()
            if i >= 2:
                best = min(best, e0.elapsed_time(e1))
        return 2.0 * n ** 3 / (best * 1e-3) / 1e12
    finally:
        torch.backends.cuda.matmul.allow_tf32 = prev


def measured_traffic(a, category):
    """DRAM bytes per launch of `category` from the committed ncu capture (profiles/ncu_traffic.json), only for the
    configuration that capture was taken on; None otherwise."""
    if not (a.kind == "copolymer" and a.batch == 4096 and a.hidden == 300 and a.depth == 3 and a.math != "fp32"):
        return None
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        return t["bytes_per_launch"].get(category)
    except Exception:
        return None


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            p = json.load(open(path))
            return float(p["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class HostGraph:
    """A fresh, uncached batch object over pinned HOST tensors (the reference's calling convention:
    get_components() returns CPU tensors, mpn.py:81-87) -- used by the end-to-end leg."""

    def __init__(self, comps):
        self._c = comps

    def get_components(self, atom_messages=False):
        return self._c


def small_batch_run(a, cpu):
    """BASELINE configs[1] also names batch 50 -- the reference's default batch size (chemprop/args.py:95).  The same
    bench in a child process at --batch 50 (the GPU is idle by now): device-resident `value`, host-buffer `e2e`, the
    forward-only figure, next to the CPU port timed at batch 50 on this host in this run."""
    cmd = [sys.executable, os.path.abspath(__file__), "--batch", "50", "--steps", "200", "--warmup", "20", "--hidden",
           str(a.hidden), "--depth", str(a.depth), "--kind", a.kind, "--math", a.math, "--no-cpu-baseline", "--no-small-batch", "--no-single-pass"]
    try:
        out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
        d = json.loads(out.stdout.strip().splitlines()[-1])
    except Exception as e:                      # never lose the main line to the side measurement
        return {"error": f"{type(e).__name__}: {e}"[:300]}
    res = {"batch": 50, "value": d["value"], "unit": UNIT, "ms_per_step": d["ms_per_step"], "gpu_launches_per_step": d["gpu_launches"] / d["steps"],
           "graph_launches": d.get("graph_launches"), "n_atoms": d["config"]["n_atoms"], "n_bonds": d["config"]["n_bonds"],
           "e2e": d["e2e"] and {k: d["e2e"][k] for k in ("value", "ms_per_step", "ms_per_step_median", "h2d_bytes_per_step", "d2h_bytes_per_step")},
           "e2e_device_store": d["e2e_device_store"] and {k: d["e2e_device_store"][k] for k in ("value", "ms_per_step")},
           "inference": d["inference"] and {k: d["inference"][k] for k in ("value", "ms_per_step")},
           "note": "each forward / backward call is ONE CUDA graph launch at this size (csrc/encoder.cu GraphRun); the "
                   "step is bound by per-kernel latency on the GPU and by Python on the host, not by HBM"}
    b50 = cpu and cpu.get("other_figures", {}).get("batch_50_all_cores")
    if b50:
        res["cpu_port_batch_50"] = {k: b50[k] for k in ("value", "threads", "ms_per_step")}
        res["value_over_cpu"] = d["value"] / b50["value"]
        res["e2e_over_cpu"] = d["e2e"]["value"] / b50["value"] if d["e2e"] else None
    return res


class EpochSampler:
    """Indices of consecutive batches of a freshly shuffled epoch (what the reference's MoleculeSampler does,
    chemprop/data/data.py:420-470: one permutation per epoch, sliced in order) -- not a new permutation per batch."""

    def __init__(self, n, batch, seed):
        self.n, self.batch, self.rng, self.perm, self.at = n, batch, np.random.default_rng(seed), None, 0

    def next(self):
        if self.perm is None or self.at + self.batch > self.n:
            self.perm, self.at = self.rng.permutation(self.n), 0
        out = self.perm[self.at:self.at + self.batch]
        self.at += self.batch
        return out


def run_train_step(a, mpn, params, store, rng, batch, dev, world, copy_stream, sync_all):
    """"a second line WITH FFN + Adam for the DP config" -- one optimisation step as
    chemprop/train/train.py:46-86 runs it: batch from the device-resident store -> encoder -> FFN head (model.py:57-116,
    2 layers, hidden 300, 1 task) -> masked MSE -> backward -> gradient all-reduce -> FlatAdam step."""
    import torch.distributed as dist
    import torch.nn as nn
    from polymer_chemprop_b200 import FlatAdam, dp, ffn_mse_loss
    torch.manual_seed(1)
    ffn = nn.Sequential(nn.Dropout(0.0), nn.Linear(a.hidden, 300), nn.ReLU(), nn.Dropout(0.0), nn.Linear(300, 1)).to(dev)
    allp = list(params) + list(ffn.parameters())
    reducer = dp.FlatGradAllReduce(allp) if world > 1 else None
    if reducer is not None:
        reducer.attach(mpn)
    opt = FlatAdam(allp, lr=1e-4, grad_buffer=reducer)
    targets = torch.randn(batch, 1, device=dev)
    mask = torch.ones(batch, 1, device=dev)
    state = {"next": store.batch(rng.next(), stream=copy_stream)}

    def one():
        cur = state["next"]
        state["next"] = store.batch(rng.next(), stream=copy_stream)
        for p in allp:
            p.grad = None
        _preds, loss = ffn_mse_loss(mpn([cur]), ffn, targets, mask)     # fused head + masked MSE (wd_head_forward / backward)
        loss.backward()
        if reducer is not None:
            reducer()
        opt.step()
        return loss

    for _ in range(10):
        one()
    sync_all()
    k = max(3, min(a.steps, 20 if batch >= 1024 else 200))     # small batches: more steps, one host stall must not be the mean
    walls = []
    gc.collect()
    gc.disable()                                  # as timeit does: no collector pause inside the timed steps
    one()                                         # one more untimed step: what the collection released is re-acquired here
    sync_all()
    t0 = time.perf_counter()
    for _ in range(k):
        ts = time.perf_counter()
        loss = one()
        walls.append(1e3 * (time.perf_counter() - ts))     # host time to ENQUEUE the step (nothing is waited for)
    float(loss.detach())                          # the step's loss reaches the host
    sync_all()
    gc.enable()
    te = torch.tensor([(time.perf_counter() - t0) / k], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    if reducer is not None:
        reducer.detach()
    return {"value": world * batch / float(te.item()), "unit": UNIT, "ms_per_step": 1e3 * float(te.item()), "steps": k,
            "host_enqueue_ms_median": float(np.median(walls)), "host_enqueue_ms": [round(w, 3) for w in walls[:40]],
            "what": "store batch -> MPN -> fused FFN(300, 1) + masked MSE (wd_head_*) -> backward -> all-reduce -> FlatAdam, "
                    "wall clock; the step's loss is read on the host at the end of the timed run"}


def main():
    a = parse_args()
    if a.impl == "reference":
        return run_reference(a)
    import torch.distributed as dist
    from polymer_chemprop_b200 import MPN, BatchMolGraph, DeviceGraph, _capi, synth

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (there is no CPU fallback for the hot path)")
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    # per-rank shard of the global batch: independent graphs, no data-path collective (weak scaling)
    graphs = synth.make_graphs(a.kind, a.batch, seed=rank)
    bmg = BatchMolGraph(graphs)
    model = MPN(make_model_args(a, dev), atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(dev)
    init_weights(model)
    params = [p for p in model.parameters() if p.requires_grad]
    dg = DeviceGraph.get(bmg, dev)                       # inputs resident in HBM before timing
    Na, Nb, G = dg.n_atoms, dg.n_bonds, dg.n_mols
    d_out = torch.randn(G, a.hidden, device=dev, generator=torch.Generator(device=dev).manual_seed(1))

    from polymer_chemprop_b200 import dp
    reducer = dp.FlatGradAllReduce(params) if world > 1 else None
    if reducer is not None:
        reducer.attach(model)      # wd_encoder_backward writes straight into the flat buffer: no pack / unpack kernels

    def step():
        for p in params:
            p.grad = None
        out = model([bmg])
        out.backward(d_out)
        if world > 1:                                    # data-parallel gradient all-reduce (one flat buffer)
            reducer()
        return out

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()                                  # polls from here on; only the timed passes are counted
    for _ in range(max(a.warmup, 3)):
        step()
    sync_all()
    lib = _capi.lib()
    lib.wd_profile_reset()

    def timed(k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sync_all()
        e0.record()
        for _ in range(k):
            step()
        e1.record()
        sync_all()
        t = torch.tensor([e0.elapsed_time(e1)], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / k

    # pass 1 -- the timed region of `value`: exactly K steps, nothing between the launches
    sampler.begin()
    n0 = _capi.launch_count()
    ms_step = timed(a.steps)
    launches = _capi.launch_count() - n0
    # pass 2 -- the same K steps with a CUDA-event pair around every launch of the library (on the launch stream),
    # for the per-kernel durations behind `roofline` / `kernels`.  The events serialise the launches (no programmatic
    # dependent launch overlap), so this pass is a little slower than pass 1: reported as ms_per_step_with_events.
    lib.wd_profile_enable(1)
    ms_step_events = timed(a.steps)
    lib.wd_profile_enable(0)
    sampler.end()
    clocks = sampler.stop() if rank == 0 else None
    prof = _capi.profile_read()
    lib.wd_profile_reset()
    value = world * G / (ms_step * 1e-3)

    # ---- share of the step spent in the gradient all-reduce (SURVEY.md 8(e) "Report"): the collective alone, back to back
    allreduce = None
    if world > 1:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sync_all()
        e0.record()
        for _ in range(a.steps):
            reducer()
        e1.record()
        sync_all()
        t = torch.tensor([e0.elapsed_time(e1) / a.steps], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        allreduce = {"ms_per_step": float(t.item()), "bytes": reducer.nbytes, "share_of_step": float(t.item()) / ms_step,
                     "note": "one ncclAllReduce(avg) of the flat gradient buffer per step, issued after the backward pass; the "
                             "encoder's backward writes its gradients straight into that buffer"}

    # ---- end to end: pinned host tensors -> H2D -> feature expansion + CSR build -> fwd+bwd -> D2H of the output
    e2e, e2e_full, e2e_store, e2e_step_sync = None, None, None, None
    # the look-ahead loop keeps two steps' workspaces alive at once (the previous step's output is still on its way to
    # the host when the next forward pass allocates): a configuration whose training workspace is over ~45 % of the
    # device memory (configs[3]: 92 GB at h=2400, d=6, B=8192) has no such loop on one GPU -- say so instead of failing
    e2e_skipped = None
    if not a.no_e2e:
        ws_train = int(lib.wd_workspace_bytes(C.byref(model.encoder[0].last_config), Na, Nb, G))
        if 2.2 * ws_train > torch.cuda.get_device_properties(dev).total_memory:
            a.no_e2e = True
            e2e_skipped = (f"training workspace {ws_train / 2**30:.1f} GiB: two steps in flight do not fit this GPU; "
                           "device-resident figures only")
    if not a.no_e2e:
        comps = bmg.get_components()
        pinned = tuple(x.pin_memory() for x in comps[:7]) + tuple(comps[7:])
        bmg.pin_memory()                                 # once, outside the timed region (DataLoader pin_memory=True)
        host_out = torch.empty(G, a.hidden).pin_memory()
        d2h = host_out.numel() * 4
        copy_stream = torch.cuda.Stream(device=dev)
        d2h_stream = torch.cuda.Stream(device=dev)

        def nbytes(ts):
            return sum(x.numel() * x.element_size() for x in ts)

        def run_e2e(native, sync="result"):
            """The public API with a one-batch look-ahead, as a data loader drives it: the upload + device-side
            graph build of batch i+1 run on a copy stream (DeviceGraph.prefetch) while the encoder step of
            batch i runs.  Every step uploads one full batch from pinned host memory and reads its encoder
            output back.  native: the batch is this package's BatchMolGraph (a fresh object per step, nothing
            cached on it), which carries f_bonds_tail and array forms of a_scope / degree_of_polym, so f_bonds is
            rebuilt on the device instead of being uploaded; otherwise a foreign object that only offers
            get_components() (Python lists for the scopes, all seven tensors uploaded as they are).
            sync: "result" -- the host blocks every step until that step's encoder output has arrived in pinned host
            memory (what a training loop's loss.item() waits for) and then launches the next step while the backward
            pass is still running; "step" -- the host also waits for the backward pass (and the gradient all-reduce)
            before it launches anything of the next step, so the GPU idles during the host's launch prologue."""
            def next_batch():
                if native == "store":                    # a freshly shuffled batch assembled in HBM from the resident store
                    return store.batch(store_rng.next(), stream=copy_stream)
                hg = bmg.fresh_view() if native else HostGraph(pinned)
                DeviceGraph.prefetch(hg, dev, copy_stream)
                return hg

            state = {"next": next_batch()}

            def e2e_step():
                cur = state["next"]
                state["next"] = next_batch()             # upload of batch i+1 overlaps the step of batch i
                for p in params:
                    p.grad = None
                out = model([cur])
                fwd_done = torch.cuda.Event()
                fwd_done.record()
                with torch.cuda.stream(d2h_stream):      # the result leaves on its own stream while the backward runs
                    d2h_stream.wait_event(fwd_done)
                    host_out.copy_(out.detach(), non_blocking=True)
                out.detach().record_stream(d2h_stream)
                out.backward(d_out)
                if world > 1:
                    reducer()
                if sync == "step":                       # host waits for the whole step (backward, all-reduce) as well
                    torch.cuda.current_stream().synchronize()
                d2h_stream.synchronize()                 # the step's result is on the host (what loss.item() waits for)

            # allocator / pinned-buffer pools reach their steady state: with 5 steps the first loop of a process still left
            # 17-33 ms of one-off work to the final synchronisation (scripts/e2e_tail_probe.py), with 15 it is gone
            for _ in range(15):
                e2e_step()
            sync_all()
            k = max(3, min(a.steps, 20 if G >= 1024 else 200))     # small batches: more steps, one host stall must not be the mean
            walls = []
            gc.collect()
            gc.disable()                             # as timeit does: no collector pause inside the timed steps
            e2e_step()                               # one more untimed step: what the collection released is re-acquired here
            sync_all()
            t0 = time.perf_counter()
            for _ in range(k):
                ts = time.perf_counter()
                e2e_step()
                walls.append(1e3 * (time.perf_counter() - ts))
            sync_all()
            gc.enable()
            te = torch.tensor([(time.perf_counter() - t0) / k], device=dev, dtype=torch.float64)
            if world > 1:
                dist.all_reduce(te, op=dist.ReduceOp.MAX)
            up = [bmg.f_atoms, bmg.f_bonds_tail, bmg.w_atoms, bmg.w_bonds, bmg.a2b, bmg.b2a, bmg.b2revb,
                  bmg.a_scope_array, bmg.dop_array] if native else list(pinned[:7])
            if native == "store":
                return {"value": world * G / float(te.item()), "unit": UNIT, "ms_per_step": 1e3 * float(te.item()),
                        "h2d_bytes_per_step": int(24 * G), "d2h_bytes_per_step": int(d2h), "steps": k,
                        "ms_per_step_median": float(np.median(walls)), "step_ms": [round(w, 3) for w in walls[:40]], "store_bytes": int(store.nbytes), "host_sync": sync,
                        "api": "DeviceGraphStore.batch(shuffled indices, copy stream) || MPN.forward(batch) + backward + "
                               "out -> pinned host; the featurised dataset is resident in HBM (uploaded once, outside the "
                               "timed region), each step uploads only the selected indices and their offsets"}
            return {"value": world * G / float(te.item()), "unit": UNIT, "ms_per_step": 1e3 * float(te.item()),
                    "h2d_bytes_per_step": int(nbytes(up) + (0 if native else G * 12)), "d2h_bytes_per_step": int(d2h),
                    "steps": k, "ms_per_step_median": float(np.median(walls)), "step_ms": [round(w, 3) for w in walls[:40]], "host_sync": sync,
                    "api": "MPN.forward(batch) + backward + out -> pinned host || DeviceGraph.prefetch(next batch of "
                           "pinned CPU tensors, copy stream)",
                    "upload": ("f_atoms + bond-only tail of f_bonds (BatchMolGraph.f_bonds_tail) + weights + indices; "
                               "f_bonds rebuilt in HBM by wd_expand_features") if native else
                              "all seven get_components() tensors as they are"}

        train_step = None
        e2e_full = run_e2e(False)
        e2e = run_e2e(True) if bmg.f_bonds_tail is not None else e2e_full
        e2e_step_sync = run_e2e(bmg.f_bonds_tail is not None, sync="step")
        e2e_store = None
        if bmg.f_bonds_tail is not None:
            from polymer_chemprop_b200 import DeviceGraphStore
            # BASELINE configs[4] / SURVEY 8(d) config 5: "dataset pre-staged on device", 10^6 graphs over 8 GPUs
            n_store = a.store_graphs or (125000 if world > 1 else len(graphs))
            store = DeviceGraphStore(graphs, dev, repeat=max(1, -(-n_store // len(graphs))))
            store_rng = EpochSampler(len(store), len(graphs), rank)
            e2e_store = run_e2e("store")
            e2e_store["store_graphs_per_rank"] = len(store)
            e2e_store["distinct_graphs_per_rank"] = len(graphs)
            train_step = run_train_step(a, model, params, store, store_rng, len(graphs), dev, world, copy_stream, sync_all)

    # ---- forward-only inference path (predict / fingerprint; wd_encoder_infer) ----------------
    inference = None
    if rank == 0:
        with torch.no_grad():
            for _ in range(3):
                model([bmg])
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            n1 = _capi.launch_count()
            e0.record()
            for _ in range(a.steps):
                model([bmg])
            e1.record()
            torch.cuda.synchronize()
        ms_inf = e0.elapsed_time(e1) / a.steps
        cfg_inf = model.encoder[0].last_config
        inference = {"value": G / (ms_inf * 1e-3), "unit": UNIT, "ms_per_step": ms_inf,
                     "gpu_launches_per_step": (_capi.launch_count() - n1) / a.steps,
                     "workspace_bytes": int(_capi.lib().wd_infer_workspace_bytes(C.byref(cfg_inf), Na, Nb, G)),
                     "training_workspace_bytes": int(_capi.lib().wd_workspace_bytes(C.byref(cfg_inf), Na, Nb, G)),
                     "note": "MPN.forward under torch.no_grad() on the resident batch of one GPU: forward kernels only, "
                             "two alternating message buffers, nothing kept for a backward pass"}

    # ---- the same step in the other tensor-core modes: single-pass TF32 (accuracy 1e-3) and the split at TF32 / bfloat16 ----
    other_modes = {}
    if rank == 0 and a.math in ("tf32x3", "bf16x3") and not a.no_single_pass:
        for other in [m for m in ("tf32", "tf32x3", "bf16x3") if m != a.math]:
            a1 = argparse.Namespace(**vars(a))
            a1.math = other
            m1 = MPN(make_model_args(a1, dev), atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(dev)
            m1.load_state_dict(model.state_dict())
            p1 = [p for p in m1.parameters() if p.requires_grad]

            def step1():
                for p in p1:
                    p.grad = None
                m1([bmg]).backward(d_out)

            for _ in range(3):
                step1()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(a.steps):
                step1()
            e1.record()
            torch.cuda.synchronize()
            ms1 = e0.elapsed_time(e1) / a.steps
            lib.wd_profile_reset()
            lib.wd_profile_enable(1)
            for _ in range(a.steps):
                step1()
            torch.cuda.synchronize()
            lib.wd_profile_enable(0)
            other_modes[other] = {"ms": ms1, "prof": _capi.profile_read()}
            lib.wd_profile_reset()
            del m1, p1

    # ---- optimiser step over the encoder's parameters (not part of `value` / `e2e`) ----------
    optimizer_step = None
    if rank == 0 and not a.no_e2e:
        from polymer_chemprop_b200 import FlatAdam

        def time_opt(make):
            ps = [torch.nn.Parameter(p.detach().clone()) for p in params]
            for q in ps:
                q.grad = torch.randn_like(q)
            opt = make(ps)
            for _ in range(5):
                opt.step()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(50):
                opt.step()
            e1.record()
            torch.cuda.synchronize()
            return 1e3 * e0.elapsed_time(e1) / 50

        optimizer_step = {"params": int(sum(p.numel() for p in params)),
                          "flat_adam_us": time_opt(lambda ps: FlatAdam(ps, lr=1e-4)),
                          "flat_adam_clip_us": time_opt(lambda ps: FlatAdam(ps, lr=1e-4, max_grad_norm=1.0)),
                          "torch_adam_us": time_opt(lambda ps: torch.optim.Adam(ps, lr=1e-4)),
                          "note": "device time per optimizer.step() over 50 back-to-back steps (host-launch bound for torch's "
                                  "for-each Adam); FlatAdam = one wd_adam_step launch (three with gradient clipping)"}

    if rank == 0:
        peak, peak_src = load_peaks()
        sign_bits = a.math != "fp32" and os.environ.get("WD_SIGN_BITS", "1") != "0"    # model activation is ReLU
        alg = algorithmic_bytes(Na, Nb, G, a.hidden, bmg.atom_fdim, bmg.bond_fdim, a.depth, sign_bits)
        kernels = {}
        gpu_ms = sum(ms for ms, _n in prof.values())
        for name, (ms, n) in prof.items():
            if n == 0:
                continue
            per = ms / n
            gbs = alg[name] / (per * 1e-3) / 1e9 if alg[name] else None
            kernels[name] = {"launches_per_step": n / a.steps, "ms_per_launch": per,
                             "share_of_gpu_time": ms / gpu_ms if gpu_ms else None,
                             "alg_bytes_per_launch": alg[name], "achieved_gbs": gbs,
                             "frac_of_hbm_peak": gbs / peak if gbs else None}
        dom = max((k for k in kernels if k not in ("misc", "reduce")), key=lambda k: kernels[k]["share_of_gpu_time"])
        roof = {"kernel": dom, "bound": "hbm", "achieved": kernels[dom]["achieved_gbs"], "peak": peak,
                "unit": "GB/s", "frac": kernels[dom]["frac_of_hbm_peak"], "traffic": measured_traffic(a, dom),
                "peak_source": peak_src, "ms_per_launch": kernels[dom]["ms_per_launch"],
                "alg_bytes_per_launch": alg[dom]}
        step_bytes = sum(alg[k] * v["launches_per_step"] for k, v in kernels.items())
        # the contract's own byte / flop counts next to the as-implemented ones
        cb = contract_bytes(Na, Nb, G, a.hidden, bmg.atom_fdim, bmg.bond_fdim, a.depth)
        cf = contract_flops(Na, Nb, a.hidden, bmg.atom_fdim, bmg.bond_fdim, a.depth)
        roof["alg_bytes_source"] = "as implemented (DESIGN.md 4.1); frac_contract uses SURVEY.md 8(d)'s bytes"
        if dom == "depth_step" and cb["depth_step"]:
            roof["contract_bytes_per_launch"] = cb["depth_step"]
            roof["frac_contract"] = cb["depth_step"] / (kernels[dom]["ms_per_launch"] * 1e-3) / 1e9 / peak
        step_tensor = None
        if cf["step"] / cb["step"] > 150.0:        # above the TF32 ridge (~100-170 flop/B): tensor-bound (BASELINE configs[3])
            tf32_peak = measure_tf32_peak(dev)
            per = kernels["depth_step"]["ms_per_launch"]
            ach = cf["depth_step"] / (per * 1e-3) / 1e12
            roof = {"kernel": "depth_step", "bound": "tensor", "achieved": ach, "peak": tf32_peak, "unit": "TFLOP/s",
                    "frac": ach / tf32_peak, "traffic": None, "ms_per_launch": per,
                    "peak_source": "measured here: torch.matmul fp32 with allow_tf32, 8192^3, best of 10 (burst), the way "
                                   "MEASURED_PEAKS.json measures bf16", "flops_per_launch": cf["depth_step"]}
            step_tensor = {"flops": cf["step"], "achieved_tflops": cf["step"] / (ms_step * 1e-3) / 1e12,
                           "frac": cf["step"] / (ms_step * 1e-3) / 1e12 / tf32_peak}
        notes = {"tf32": "same kernels with one tcgen05.mma kind::tf32 pass per K step (encoder output within 1e-3 of the "
                         "float64 oracle; predictions miss the 1e-4 bound)",
                 "tf32x3": "forward contractions split at TF32: tf32 hi products + one bf16 group over the correction images "
                           "(encoder output 2e-6)",
                 "bf16x3": "forward contractions split at bfloat16: three kind::f16 products over one image per operand "
                           "(encoder output 4e-6)"}
        for other, r in other_modes.items():
            ms1, prof1 = r["ms"], r["prof"]
            o = {"math_mode": other, "value": G / (ms1 * 1e-3), "unit": UNIT, "ms_per_step": ms1, "kernels": {},
                 "note": notes[other] + "; the difference to `value` is in the three forward contractions"}
            for name in ("input_proj", "depth_step", "atom_update"):
                ms, n = prof1.get(name, (0.0, 0))
                if n:
                    o["kernels"][name] = {"ms_per_launch": ms / n, "frac_of_hbm_peak": alg[name] / (ms / n * 1e-3) / 1e9 / peak}
            other_modes[other] = o
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps,
                "warmup": max(a.warmup, 3), "ms_per_step": ms_step, "ms_per_step_with_events": ms_step_events,
                "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None,
                "dtype": {"fp32": "f32", "tf32": "tf32", "tf32x3": "tf32x3", "bf16x3": "bf16x3"}[a.math],
                "data": "synthetic",
                "config": base_config(a, world, G, Na, Nb),
                "options": {"math_mode": a.math, "pdl": os.environ.get("WD_PDL", "1") != "0", "sign_bits": sign_bits,
                            "e2e_skipped": e2e_skipped},
                "edge_msgs_per_s": world * (Nb - 1) * max(a.depth - 1, 0) / (ms_step * 1e-3),
                "roofline": roof, "kernels": kernels,
                "step_hbm": {"alg_bytes": step_bytes, "achieved_gbs": step_bytes / (ms_step * 1e-3) / 1e9,
                             "frac": step_bytes / (ms_step * 1e-3) / 1e9 / peak,
                             "contract_bytes": cb["step"], "frac_contract": cb["step"] / (ms_step * 1e-3) / 1e9 / peak},
                "step_tensor": step_tensor,
                "gpu_launches": int(launches), "clocks": clocks, "e2e": e2e, "e2e_full_tensors": e2e_full,
                "e2e_device_store": e2e_store, "e2e_host_waits_for_backward": e2e_step_sync,
                "optimizer_step": optimizer_step, "inference": inference, "allreduce": allreduce,
                "train_step": train_step if not a.no_e2e else None, "single_pass_tf32": other_modes.get("tf32"),
                "other_modes": {k: v for k, v in other_modes.items() if k != "tf32"} or None}
        if world > 1 and e2e_store is not None:
            # data-parallel configuration = BASELINE configs[4]: the dataset is pre-staged on the device, so the end-to-end
            # figure is the store-fed one; the host-upload loop stays alongside
            line["e2e_host_upload"] = e2e
            line["e2e"] = e2e_store
        if not a.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(graphs, a, steps=10, warmup=2, extras=True)
        line["graph_launches"] = dict(zip(("launches", "executable_graphs_built"), _capi.graph_stats()))
        if world == 1 and a.batch > 50 and not a.no_small_batch and not a.no_e2e:
            line["small_batch"] = small_batch_run(a, line.get("cpu_baseline"))
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
