"""Per-role wait-cycle breakdown of the tcgen05 GEMM kernel (library built with -DWD_T2_TIMING)."""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from polymer_chemprop_b200 import _capi  # noqa: E402

DEV = torch.device("cuda", 0)
L, p = _capi.lib(), _capi.ptr
NAMES = ["xform<-raw_full", "xform<-a_empty", "loader<-raw_empty", "bload<-b_empty", "mma<-tmem_empty", "mma<-a_full",
         "mma<-b_full", "epi<-tmem_full", "epi<-r_full", "rload<-r_empty"]


def run(M, N, K, epi):
    A = torch.randn(M, K, device=DEV)
    W = torch.randn(N, K, device=DEV) / K ** 0.5
    R = torch.randn(M, (N + 7) // 8 * 8, device=DEV) if epi == 2 else None
    ldc = (N + 7) // 8 * 8
    Cc = torch.empty(M, ldc, device=DEV)
    scratch = torch.empty(4 * (N + 320) * (K + 64) + 1024, dtype=torch.uint8, device=DEV)
    buf = (C.c_uint64 * 16)()
    for it in range(3):
        L.wd_debug_timing(buf, 1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _capi.check(L.wd_debug_gemm(p(A), A.stride(0), p(W), W.stride(0), 0, None, p(R), R.stride(0) if R is not None else 0,
                                    p(Cc), ldc, M, N, K, epi, 1, p(scratch), scratch.numel(), _capi.current_stream(DEV)))
        e1.record()
        torch.cuda.synchronize()
    L.wd_debug_timing(buf, 0)
    tot = buf[15] or 1
    print(f"M={M} N={N} K={K} epi={epi}: {e0.elapsed_time(e1) * 1e3:.0f} us (incl. pack), CTA0 cycles {tot}")
    for i, n in enumerate(NAMES):
        print(f"   {n:20s} {100.0 * buf[i] / tot:6.1f}% of kernel")


def run_dw(Rn, N1, N2):
    X = torch.randn(Rn, (N1 + 7) // 8 * 8, device=DEV)
    Y = torch.randn(Rn, (N2 + 3) // 4 * 4, device=DEV)
    out = torch.empty(N1, N2, device=DEV)
    scratch = torch.empty(160 * N1 * N2 * 4, dtype=torch.uint8, device=DEV)
    buf = (C.c_uint64 * 16)()
    for it in range(3):
        L.wd_debug_timing(buf, 1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _capi.check(L.wd_debug_dw(p(X), X.stride(0), p(Y), Y.stride(0), p(out), Rn, N1, N2, 1, p(scratch), scratch.numel(),
                                  _capi.current_stream(DEV)))
        e1.record()
        torch.cuda.synchronize()
    L.wd_debug_timing(buf, 0)
    tot = buf[15] or 1
    print(f"dW R={Rn} N1={N1} N2={N2}: {e0.elapsed_time(e1) * 1e3:.0f} us (incl. reduce), CTA0 cycles {tot}")
    for i, n in ((0, "xform<-raw_full"), (2, "loader<-empty"), (3, "tma<-empty"), (5, "mma<-full"), (6, "mma<-op_full"),
                 (7, "epi<-tmem_full")):
        print(f"   {n:20s} {100.0 * buf[i] / tot:6.1f}% of kernel")


def run_encoder():
    """Wait-cycle counters of whatever instrumented kernels one encoder forward + backward launches (build the
    timing library with -DWD_T2_TIMING_ONLY=n to isolate one kernel generation, e.g. 4 = gathered depth step)."""
    import types
    from polymer_chemprop_b200 import MPN, BatchMolGraph, synth
    bmg = BatchMolGraph(synth.make_graphs("copolymer", 4096, seed=0))
    args = types.SimpleNamespace(
        hidden_size=300, depth=3, bias=False, dropout=0.0, activation="ReLU", atom_messages=False, undirected=False,
        aggregation="mean", aggregation_norm=100, atom_descriptors=None, atom_descriptors_size=0, device=DEV,
        mpn_shared=False, number_of_molecules=1, features_only=False, use_input_features=False,
        overwrite_default_atom_features=False, overwrite_default_bond_features=False, wd_math_mode="tf32")
    model = MPN(args, atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(DEV)
    buf = (C.c_uint64 * 16)()
    for it in range(2):
        L.wd_debug_timing(buf, 1)
        out = model([bmg])
        out.sum().backward()
        torch.cuda.synchronize()
    L.wd_debug_timing(buf, 0)
    tot = buf[15] or 1
    print(f"encoder fwd+bwd, CTA0 cycles of the instrumented kernels {tot}")
    for i, n in enumerate(NAMES):
        print(f"   {n:20s} {100.0 * buf[i] / tot:6.1f}% of kernel")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "encoder":
        run_encoder()
        sys.exit(0)
    run_dw(269541, 300, 300)
    run_dw(269541, 300, 147)
    run(269541, 300, 300, 0)
    run(269541, 300, 300, 2)
    run(269541, 300, 147, 0)
