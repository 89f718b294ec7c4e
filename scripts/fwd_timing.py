"""Per-role wait-cycle breakdown of the forward kernel k_tc_gemm4 (CTA 0) for the three forward contractions at the
benchmarked sizes, single-pass TF32 vs 3xTF32.  Needs a library built with -DWD_T2_TIMING -DWD_T2_TIMING_ONLY=4
(WDMPNN_LIB=.../libwdmpnn_timing.so)."""
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
DEV = torch.device("cuda", 0)
L, p = _capi.lib(), _capi.ptr
buf = (C.c_uint64 * 16)()


def report(title):
    L.wd_debug_timing(buf, 0)
    tot = buf[15] or 1
    print(f"{title}: CTA 0 ran {tot} cycles")
    for i, n in NAMES.items():
        print(f"   {n:36s} {100.0 * buf[i] / tot:6.1f} % of the kernel's cycles waiting")


def gemm(M, N, K, lda, epi, mode):
    A = torch.randn(M, lda, device=DEV)
    W = torch.randn(N, K, device=DEV) / K ** 0.5
    ldc = (N + 7) // 8 * 8
    R = torch.randn(M, ldc, device=DEV) if epi == 2 else None
    Cc = torch.empty(M, ldc, device=DEV)
    scratch = torch.empty(8 * (N + 320) * (K + 64) + 1024, dtype=torch.uint8, device=DEV)
    for it in range(3):
        L.wd_debug_timing(buf, 1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _capi.check(L.wd_debug_gemm(p(A), lda, p(W), K, 0, None, p(R), ldc if R is not None else 0, p(Cc), ldc, M, N, K, epi,
                                    mode, p(scratch), scratch.numel(), _capi.current_stream(DEV)))
        e1.record()
        torch.cuda.synchronize()
    report(f"plain GEMM M={M} N={N} K={K} epi={epi} mode={mode}: {e0.elapsed_time(e1) * 1e3:.0f} us incl. pack")


def depth_step(mode, batch=4096, hidden=300, depth=3):
    a = types.SimpleNamespace(hidden=hidden, depth=depth, math=mode)
    bmg = BatchMolGraph(synth.make_graphs("copolymer", batch, seed=0))
    model = MPN(bench.make_model_args(a, DEV), atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(DEV)
    bench.init_weights(model)
    with torch.no_grad():
        for it in range(3):
            L.wd_debug_timing(buf, 1)
            model([bmg])
            torch.cuda.synchronize()
    report(f"encoder forward h={hidden} d={depth} B={batch}, mode {mode} (every k_tc_gemm4 launch of one forward pass summed)")


if __name__ == "__main__":
    if sys.argv[1:2] == ["wide"]:               # configs[3] arithmetic (h = 2400, d = 6) at 2048 graphs
        for mode in ("bf16x3", "tf32x3", "tf32"):
            depth_step(mode, 2048, 2400, 6)
        sys.exit(0)
    M = 269541
    for mode in (3, 2, 1):
        gemm(M, 300, 147, 148, 0, mode)          # input projection (mode 1 runs k_tc_gemm3 unless rows are unaligned)
        gemm(M, 300, 300, 304, 2, mode)          # residual epilogue through k_tc_gemm4, plain operand
    for mode in ("bf16x3", "tf32x3", "tf32"):
        depth_step(mode)
