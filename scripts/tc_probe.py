"""Diagnostic probe of the tcgen05 contraction kernels through the C-ABI test hooks (run on the
GPU box).  Prints relative errors against an fp64 torch product; on a mismatch prints the error
structure (which rows / columns are wrong) so layout / descriptor bugs can be read off one run."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from polymer_chemprop_b200 import _capi  # noqa: E402

DEV = torch.device("cuda", 0)
L, p = _capi.lib(), _capi.ptr


def rel(a, b):
    return float((a.double() - b.double()).norm() / max(b.double().norm(), 1e-30))


def gemm(A, W, b_kn=False, bias=None, R=None, epi=0, mode=1, N=None, K=None):
    M = A.shape[0]
    K = K or A.shape[1]
    N = N or (W.shape[1] if b_kn else W.shape[0])
    ldc = (N + 3) // 4 * 4
    C = torch.full((M, ldc), float("nan"), device=DEV)
    scratch = torch.empty(4 * (N + 320) * (K + 64) + 1024, dtype=torch.uint8, device=DEV)
    rc = L.wd_debug_gemm(p(A), A.stride(0), p(W), W.stride(0), int(b_kn), p(bias), p(R), R.stride(0) if R is not None else 0,
                         p(C), ldc, M, N, K, epi, mode, p(scratch), scratch.numel(), _capi.current_stream(DEV))
    _capi.check(rc)
    torch.cuda.synchronize()
    return C


def ref_gemm(A, W, b_kn, bias, R, epi, N, K):
    Wm = (W[:K, :N].double().t() if b_kn else W[:N, :K].double())
    out = A[:, :K].double() @ Wm.t()
    if bias is not None:
        out = out + bias.double()
    if epi == 2:
        out = out + R[:, :N].double()
    if epi:
        out = out.clamp_min(0)
    return out


def probe_gemm(M, N, K, b_kn=False, epi=0, bias=False, lda=None):
    g = torch.Generator(device=DEV).manual_seed(M * 7 + N * 3 + K)
    lda = lda or K
    A = torch.randn(M, lda, device=DEV, generator=g)
    W = torch.randn((K, N) if b_kn else (N, K), device=DEV, generator=g) / K ** 0.5
    b = torch.randn(N, device=DEV, generator=g) if bias else None
    R = torch.randn(M, (N + 3) // 4 * 4, device=DEV, generator=g) if epi == 2 else None
    ref = ref_gemm(A, W, b_kn, b, R, epi, N, K)
    out = {}
    for mode, name in ((0, "fp32"), (1, "tf32")):
        try:
            C = gemm(A, W, b_kn, b, R, epi, mode, N, K)
        except Exception as e:  # noqa: BLE001
            print(f"  gemm M={M} N={N} K={K} {name}: EXCEPTION {e}")
            return False
        out[name] = rel(C[:, :N], ref)
        pad_ok = bool((C[:, N:] == 0).all())
        if name == "tf32" and not (out[name] < 2e-3):
            err = (C[:, :N].double() - ref).abs()
            bad_rows = (err.amax(1) > 1e-2).nonzero().flatten()
            bad_cols = (err.amax(0) > 1e-2).nonzero().flatten()
            print(f"    BAD tf32: nan={int(torch.isnan(C).sum())} bad_rows={bad_rows.numel()}/{M} "
                  f"(first {bad_rows[:12].tolist()}) bad_cols={bad_cols.numel()}/{N} (first {bad_cols[:12].tolist()})")
            print("    C[0,:8] =", C[0, :8].tolist())
            print("    ref[0,:8]=", ref[0, :8].float().tolist())
    print(f"  gemm M={M:6d} N={N:4d} K={K:4d} b_kn={int(b_kn)} epi={epi} bias={int(bias)} lda={lda}: "
          f"fp32 {out['fp32']:.2e}  tf32 {out['tf32']:.2e}  pad_zero={pad_ok}")
    return out["tf32"] < 2e-3 and out["fp32"] < 1e-5


def probe_dw(Rn, N1, N2):
    g = torch.Generator(device=DEV).manual_seed(Rn + N1 + N2)
    ld1, ld2 = (N1 + 3) // 4 * 4, N2
    X = torch.randn(Rn, ld1, device=DEV, generator=g)
    Y = torch.randn(Rn, ld2, device=DEV, generator=g)
    ref = X[:, :N1].double().t() @ Y[:, :N2].double()
    res = {}
    for mode, name in ((0, "fp32"), (1, "tf32")):
        out = torch.full((N1, N2), float("nan"), device=DEV)
        scratch = torch.empty(160 * N1 * N2 * 4, dtype=torch.uint8, device=DEV)
        try:
            _capi.check(L.wd_debug_dw(p(X), ld1, p(Y), ld2, p(out), Rn, N1, N2, mode, p(scratch), scratch.numel(),
                                      _capi.current_stream(DEV)))
            torch.cuda.synchronize()
        except Exception as e:  # noqa: BLE001
            print(f"  dw R={Rn} N1={N1} N2={N2} {name}: EXCEPTION {e}")
            return False
        res[name] = rel(out, ref)
        if name == "tf32" and not (res[name] < 2e-3):
            err = (out.double() - ref).abs() / ref.abs().mean()
            print(f"    BAD tf32 dw: nan={int(torch.isnan(out).sum())} bad_rows={(err.amax(1) > .05).sum().item()}/{N1} "
                  f"bad_cols={(err.amax(0) > .05).sum().item()}/{N2}")
            print("    out[0,:6]=", out[0, :6].tolist(), " ref[0,:6]=", ref[0, :6].float().tolist())
    print(f"  dw   R={Rn:7d} N1={N1:4d} N2={N2:4d}: fp32 {res['fp32']:.2e}  tf32 {res['tf32']:.2e}")
    return res["tf32"] < 2e-3 and res["fp32"] < 1e-5


def main():
    print("device:", torch.cuda.get_device_name(0))
    ok = True
    for args in [(128, 16, 32), (128, 64, 32), (128, 64, 64), (256, 64, 96), (100, 48, 40), (1000, 300, 300),
                 (1000, 300, 147), (3000, 301, 433), (20000, 300, 300), (700, 640, 200), (4000, 2400, 512)]:
        ok &= probe_gemm(*args)
    ok &= probe_gemm(1000, 300, 147, lda=147)
    ok &= probe_gemm(1000, 300, 300, b_kn=True)
    ok &= probe_gemm(1000, 300, 300, epi=2, bias=True)
    ok &= probe_gemm(1000, 300, 300, epi=1, bias=True)
    for args in [(1000, 64, 64), (4096, 128, 32), (10000, 300, 300), (5000, 300, 147), (50000, 300, 433), (3000, 640, 700)]:
        ok &= probe_dw(*args)
    print("PROBE", "OK" if ok else "FAILED")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
