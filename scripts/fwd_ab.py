"""A/B timing of the forward kernel variants (WDMPNN_LIB selects the library build): the plain 3xTF32 contraction at the
input-projection shape and the encoder forward / step in both tensor-core modes."""
import os
import sys
import types

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from polymer_chemprop_b200 import MPN, BatchMolGraph, _capi, synth  # noqa: E402

DEV = torch.device("cuda", 0)
L, p = _capi.lib(), _capi.ptr


def timed(fn, n=10):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


def gemm(M, N, K, lda, epi, mode):
    A = torch.randn(M, lda, device=DEV)
    W = torch.randn(N, K, device=DEV) / K ** 0.5
    ldc = (N + 7) // 8 * 8
    R = torch.randn(M, ldc, device=DEV) if epi == 2 else None
    Cc = torch.empty(M, ldc, device=DEV)
    scratch = torch.empty(8 * (N + 320) * (K + 64) + 1024, dtype=torch.uint8, device=DEV)
    s = _capi.current_stream(DEV)
    return timed(lambda: _capi.check(L.wd_debug_gemm(p(A), lda, p(W), K, 0, None, p(R), ldc if R is not None else 0, p(Cc), ldc,
                                                     M, N, K, epi, mode, p(scratch), scratch.numel(), s)))


def encoder(mode, bmg):
    a = types.SimpleNamespace(hidden=300, depth=3, math=mode)
    model = MPN(bench.make_model_args(a, DEV), atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(DEV)
    bench.init_weights(model)
    d_out = torch.randn(len(bmg.a_scope), 300, device=DEV)

    def fwd():
        with torch.no_grad():
            model([bmg])

    def step():
        for q in model.parameters():
            q.grad = None
        model([bmg]).backward(d_out)
    return timed(fwd), timed(step)


if __name__ == "__main__":
    tag = os.path.basename(os.environ.get("WDMPNN_LIB", "libwdmpnn.so"))
    M = 269541
    res = [f"{tag}:"]
    for mode in (2, 1, 3):
        res.append(f"gemm K=147 store m{mode} {gemm(M, 300, 147, 148, 0, mode):.0f}us")
        res.append(f"gemm K=300 res m{mode} {gemm(M, 300, 300, 304, 2, mode):.0f}us")
    bmg = BatchMolGraph(synth.make_graphs("copolymer", 4096, seed=0))
    for mode in ("tf32x3", "tf32", "bf16x3"):
        f, s = encoder(mode, bmg)
        res.append(f"{mode} fwd {f:.0f}us step {s:.0f}us")
    print(" | ".join(res), flush=True)
