"""GPU parity of the tcgen05 tensor-core paths: the raw contractions against fp64 torch products, and the
encoder against the reference's golden vectors / the CPU oracle.

Two tensor-core math modes (include/wdmpnn.h), tolerances in tests/_util.py TOL:
* tf32x3 (the default of the drop-in): error-compensated forward contractions (a_hi w_hi + a_lo w_hi + a_hi w_lo in
  one TMEM accumulator) -- encoder output <= 1e-5 relative like the FP32 path, raw contraction <= 5e-6; the backward
  contractions are single-pass TF32 on exact activations: gradients <= 5e-3 on the tiny golden batches.
* bf16x3: the same three products with the split point at bfloat16, every MMA kind::f16 (csrc/tc_gemm4.cuh): encoder
  output <= 2e-5, raw contraction <= 2e-5, gradients as tf32x3.
* tf32 (single pass): BASELINE.json's <= 1e-3 relative on the encoder output (||x-ref||_F / ||ref||_F).  Gradients
  are held to 3e-2: an exact-accumulation TF32 model of the reference (oracle/mpn_explicit.py mm_tf32,
  tests/test_oracle_golden.py::test_tf32_error_model) already shows up to 2e-2 on these tiny batches, because 3e-4
  forward perturbations flip ReLU masks."""
import numpy as np
import pytest
import torch

import _golden
import _util
from oracle import mpn_oracle
from polymer_chemprop_b200 import BatchMolGraph, _capi, synth

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)
MODES = ["tf32", "tf32x3", "bf16x3"]
MODE_ID = {"tf32": 1, "tf32x3": 2, "bf16x3": 3}
TOL = _util.TOL


def _gemm(A, W, N, K, b_kn=False, bias=None, R=None, epi=0, mode=1):
    L, p = _capi.lib(), _capi.ptr
    M, ldc = A.shape[0], (N + 3) // 4 * 4
    C = torch.full((M, ldc), float("nan"), device=DEV)
    scratch = torch.empty(8 * (N + 320) * (K + 64) + 1024, dtype=torch.uint8, device=DEV)   # hi + correction images
    _capi.check(L.wd_debug_gemm(p(A), A.stride(0), p(W), W.stride(0), int(b_kn), p(bias), p(R),
                                R.stride(0) if R is not None else 0, p(C), ldc, M, N, K, epi, mode, p(scratch),
                                scratch.numel(), _capi.current_stream(DEV)))
    torch.cuda.synchronize()
    return C


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("M,N,K", [(128, 16, 32), (100, 48, 40), (256, 64, 96), (1000, 300, 300), (1000, 300, 147),
                                   (3000, 301, 433), (20000, 300, 300), (700, 640, 200), (2000, 2400, 512)])
def test_tc_gemm_matches_fp64(M, N, K, mode):
    g = torch.Generator(device=DEV).manual_seed(M + N + K)
    A = torch.randn(M, K, device=DEV, generator=g)
    W = torch.randn(N, K, device=DEV, generator=g) / K ** 0.5
    C = _gemm(A, W, N, K, mode=MODE_ID[mode])
    ref = A.double() @ W.double().t()
    err = _golden.rel_err(C[:, :N].cpu(), ref.cpu())
    _util.record("gemm_vs_fp64", mode=mode, M=M, N=N, K=K, rel=err, max=_golden.max_err(C[:, :N].cpu(), ref.cpu()))
    assert err <= TOL[mode]["gemm"]
    assert _golden.max_err(C[:, :N].cpu(), ref.cpu()) <= 20 * TOL[mode]["gemm"]
    assert bool((C[:, N:] == 0).all())            # pad columns [N, ldc) are kept at zero


def test_tc_gemm_unaligned_rows_take_the_4_byte_copy_path():
    """Dense 133 / 147-float feature rows (WdGraph.ld_atoms / ld_bonds == 0) are only 4-byte aligned: no TMA box;
    k_tc_gemm4's loader warps copy them with 4-byte cp.async into the same swizzled stage."""
    g = torch.Generator(device=DEV).manual_seed(11)
    M, N, K = 700, 300, 147
    A = torch.randn(M, K, device=DEV, generator=g)
    W = torch.randn(N, K, device=DEV, generator=g) / K ** 0.5
    assert A.stride(0) % 4 != 0
    ref = (A.double() @ W.double().t()).cpu()
    for mode in MODES:
        C = _gemm(A, W, N, K, mode=MODE_ID[mode])
        assert _golden.rel_err(C[:, :N].cpu(), ref) <= TOL[mode]["gemm"], mode


@pytest.mark.parametrize("mode", MODES)
def test_tc_gemm_transposed_weights_and_epilogues(mode):
    tol = TOL[mode]["gemm"]
    g = torch.Generator(device=DEV).manual_seed(3)
    M, N, K = 1500, 300, 300
    A = torch.randn(M, K, device=DEV, generator=g)
    W = torch.randn(K, N, device=DEV, generator=g) / K ** 0.5
    C = _gemm(A, W, N, K, b_kn=True, mode=MODE_ID[mode])                   # backward form: single pass in both modes
    assert _golden.rel_err(C[:, :N].cpu(), (A.double() @ W.double()).cpu()) <= TOL["tf32"]["gemm"]
    Wn = torch.randn(N, K, device=DEV, generator=g) / K ** 0.5
    b = torch.randn(N, device=DEV, generator=g)
    R = torch.randn(M, N, device=DEV, generator=g)
    ref = A.double() @ Wn.double().t() + b.double()
    C1 = _gemm(A, Wn, N, K, bias=b, epi=1, mode=MODE_ID[mode])
    assert _golden.rel_err(C1[:, :N].cpu(), ref.clamp_min(0).cpu()) <= tol
    C2 = _gemm(A, Wn, N, K, bias=b, R=R, epi=2, mode=MODE_ID[mode])
    assert _golden.rel_err(C2[:, :N].cpu(), (ref + R.double()).clamp_min(0).cpu()) <= tol


@pytest.mark.parametrize("R,N1,N2", [(1000, 64, 64), (4100, 128, 32), (10000, 300, 300), (5000, 300, 147),
                                     (50000, 300, 433), (3000, 640, 700)])
def test_tc_dw_matches_fp64(R, N1, N2):
    L, p = _capi.lib(), _capi.ptr
    g = torch.Generator(device=DEV).manual_seed(R + N1 + N2)
    ld1 = (N1 + 3) // 4 * 4
    X = torch.randn(R, ld1, device=DEV, generator=g)
    Y = torch.randn(R, N2, device=DEV, generator=g)
    out = torch.full((N1, N2), float("nan"), device=DEV)
    scratch = torch.empty(160 * N1 * N2 * 4, dtype=torch.uint8, device=DEV)
    _capi.check(L.wd_debug_dw(p(X), ld1, p(Y), N2, p(out), R, N1, N2, 1, p(scratch), scratch.numel(),
                              _capi.current_stream(DEV)))
    torch.cuda.synchronize()
    ref = X[:, :N1].double().t() @ Y.double()
    assert _golden.rel_err(out.cpu(), ref.cpu()) <= TOL["tf32"]["gemm"]


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("name", _golden.NAMES)
def test_encoder_tf32_matches_reference_golden(name, mode):
    g = _golden.load(name)
    model = _util.model_from_params(g["params"], g["cfg"], math_mode=mode)
    out, grads = _util.run_fwd_bwd(model, _util.ComponentsGraph(g["components"]), g["d_out"])
    assert out.shape == g["out"].shape
    gerr = {k: _golden.rel_err(grads[k], ref) for k, ref in g["grads"].items()}
    _util.record("golden", mode=mode, name=name, out=_golden.rel_err(out, g["out"]), out_max=_golden.max_err(out, g["out"]),
                 grad=max(gerr.values()))
    assert _golden.rel_err(out, g["out"]) <= TOL[mode]["out"]
    assert _golden.max_err(out, g["out"]) <= 20 * TOL[mode]["out"]
    for k, e in gerr.items():
        assert e <= TOL[mode]["grad"], k


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("hidden,depth,undirected,bias", [(76, 4, False, False), (301, 3, True, True), (640, 2, False, True)])
def test_encoder_tf32_matches_oracle_synthetic(hidden, depth, undirected, bias, mode):
    graphs = synth.make_graphs("copolymer", 40, 21, dense_features=True, random_edge_weights=True)
    bmg = BatchMolGraph(graphs + [synth.empty_graph()])
    cfg = dict(hidden_size=hidden, depth=depth, undirected=undirected, bias=bias)
    torch.manual_seed(5)
    model = _util.MPN(_util.make_args(wd_math_mode=mode, **cfg), atom_fdim=bmg.atom_fdim,
                      bond_fdim=bmg.bond_fdim).to(DEV)
    for p in model.parameters():
        if p.requires_grad:
            torch.nn.init.normal_(p, 0.0, 0.1) if p.dim() == 1 else torch.nn.init.xavier_normal_(p)
    d_out = torch.randn(len(bmg.a_scope), hidden, generator=torch.Generator().manual_seed(9))
    out, grads = _util.run_fwd_bwd(model, bmg, d_out)
    weights = {k: v.detach().cpu() for k, v in model.encoder[0].state_dict().items() if k != "cached_zero_vector"}
    ref_out, ref_grads = mpn_oracle.encoder_forward_backward(bmg.get_components(), weights, d_out, depth=depth,
                                                             undirected=undirected)
    _util.record("synthetic", mode=mode, hidden=hidden, out=_golden.rel_err(out, ref_out),
                 grad=max(_golden.rel_err(grads[k], ref) for k, ref in ref_grads.items()))
    assert _golden.rel_err(out, ref_out) <= TOL[mode]["out"]
    for k, ref in ref_grads.items():
        assert _golden.rel_err(grads[k], ref) <= TOL[mode]["grad"], k


@pytest.mark.parametrize("mode", MODES)
def test_tf32_backward_is_bitwise_deterministic(mode):
    g = _golden.load("copoly_b50")
    model = _util.model_from_params(g["params"], g["cfg"], math_mode=mode)
    graph = _util.ComponentsGraph(g["components"])
    out1, grads1 = _util.run_fwd_bwd(model, graph, g["d_out"])
    out2, grads2 = _util.run_fwd_bwd(model, graph, g["d_out"])
    assert torch.equal(out1, out2)
    for k in grads1:
        assert torch.equal(grads1[k], grads2[k]), k


@pytest.mark.parametrize("hidden,depth,p", [(300, 3, 0.2), (96, 4, 0.4)])
@pytest.mark.parametrize("mode", MODES)
def test_dropout_tf32_matches_oracle_with_the_same_philox_masks(hidden, depth, p, mode):
    """Dropout in the tcgen05 epilogues (k_tc_gemm4 depth step and, in the tf32x3 mode, atom update; k_tc_gemm3 atom
    update in the single-pass mode)."""
    _util.dropout_case(mode, hidden, depth, "ReLU", p, TOL[mode]["out"], TOL[mode]["grad"])


def _big_batch():
    return BatchMolGraph(synth.make_graphs("copolymer", 300, 61, random_edge_weights=True) + [synth.empty_graph()] +
                         synth.make_graphs("qm9", 50, 62))


def _model(bmg, hidden=300, depth=4, mode="tf32", **kw):
    torch.manual_seed(1)
    model = _util.MPN(_util.make_args(hidden_size=hidden, depth=depth, wd_math_mode=mode, **kw),
                      atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(DEV)
    for p in model.parameters():
        if p.requires_grad and p.dim() == 2:
            torch.nn.init.xavier_normal_(p)
    return model


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("undirected", [False, True])
def test_reverse_message_tma_box_equals_gathered_rows_bitwise(undirected, mode):
    """Gathered depth step: message[b2revb] fetched as one TMA box per tile (reverse bonds are adjacent pairs in
    every reference-built graph, featurization.py:470-480 / :616-629; WdGraph.rev_adjacent) against the gathered
    cp.async rows -- same arithmetic, so the whole forward and backward must agree bit for bit."""
    from polymer_chemprop_b200 import DeviceGraph
    bmg = _big_batch()
    model = _model(bmg, undirected=undirected, mode=mode)
    d_out = torch.randn(len(bmg.a_scope), 300, generator=torch.Generator().manual_seed(2))
    out_tma, grads_tma = _util.run_fwd_bwd(model, bmg, d_out)
    assert DeviceGraph.get(bmg, DEV).rev_adjacent and DeviceGraph.get(bmg, DEV).struct.rev_adjacent == 1
    other = bmg.fresh_view()
    dg = DeviceGraph.get(other, DEV)
    dg.check()
    dg.struct.rev_adjacent = 0                                   # force the gathered path
    out_g, grads_g = _util.run_fwd_bwd(model, other, d_out)
    assert torch.equal(out_tma, out_g)
    for k in grads_tma:
        assert torch.equal(grads_tma[k], grads_g[k]), k


@pytest.mark.parametrize("mode", MODES)
def test_renumbered_bonds_take_the_gathered_path_and_match_the_oracle(mode):
    """A valid BatchMolGraph whose bonds are renumbered at random (reverse pairs no longer adjacent): wd_csr_build
    reports it (flags[2] > 0), the depth step gathers message[b2revb] row by row, and the result matches the oracle
    on the renumbered tensors -- and the encoder output of the original numbering (the encoder is invariant under
    a renumbering of the bonds that keeps every a2b row's order)."""
    from polymer_chemprop_b200 import DeviceGraph
    bmg = BatchMolGraph(synth.make_graphs("copolymer", 40, 71, dense_features=True, random_edge_weights=True))
    fa, fb, wa, wb, a2b, b2a, b2revb, a_scope, b_scope, dop = bmg.get_components()
    n_b = fb.shape[0]
    rng = np.random.default_rng(5)
    perm = torch.from_numpy(np.concatenate([[0], 1 + rng.permutation(n_b - 1)]).astype(np.int64))   # old id -> new id
    inv = torch.empty_like(perm)
    inv[perm] = torch.arange(n_b)
    comps = (fa, fb[inv], wa, wb[inv], perm[a2b], b2a[inv], perm[b2revb[inv]], a_scope, b_scope, dop)
    graph = _util.ComponentsGraph(comps)
    model = _model(bmg, hidden=76, depth=4, mode=mode)
    d_out = torch.randn(len(a_scope), 76, generator=torch.Generator().manual_seed(9))
    out, grads = _util.run_fwd_bwd(model, graph, d_out)
    assert not DeviceGraph.get(graph, DEV).rev_adjacent
    weights = {k: v.detach().cpu() for k, v in model.encoder[0].state_dict().items() if k != "cached_zero_vector"}
    ref_out, ref_grads = mpn_oracle.encoder_forward_backward(comps, weights, d_out, depth=4)
    assert _golden.rel_err(out, ref_out) <= TOL[mode]["out"]
    for k, ref in ref_grads.items():
        assert _golden.rel_err(grads[k], ref) <= TOL[mode]["grad"], k
    out0, _ = _util.run_fwd_bwd(model, bmg, d_out)
    assert _golden.rel_err(out, out0) <= 1e-6                    # per-row arithmetic is the same, only row order moved


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("hidden,depth,undirected,p", [(300, 4, False, 0.0), (300, 3, True, 0.0), (76, 4, False, 0.0),
                                                       (640, 2, False, 0.0), (300, 3, False, 0.25), (96, 3, False, 0.4)])
def test_sign_bits_give_the_same_gradients_as_the_stored_activations_bitwise(hidden, depth, undirected, p, mode, monkeypatch):
    """ReLU: the message gradient reads one sign bit per element (left by the tcgen05 epilogues, common.cuh
    SignBits) instead of the stored activation.  Same arithmetic (act' is 0 or 1, times the dropout scale), so
    every gradient must agree bit for bit with the path that reads the activations (WD_SIGN_BITS=0, i.e.
    WdConfig.flags & WD_FLAG_NO_SIGN_BITS: part of the call's config, so forward and backward cannot disagree).  The
    widths cover one and two accumulator halves and several column tiles per producer."""
    bmg = _big_batch()
    model = _model(bmg, hidden=hidden, depth=depth, undirected=undirected, dropout=p, mode=mode)
    model.train()
    d_out = torch.randn(len(bmg.a_scope), hidden, generator=torch.Generator().manual_seed(2))
    enc = model.encoder[0]
    res = {}
    for bits in ("1", "0"):
        monkeypatch.setenv("WD_SIGN_BITS", bits)
        torch.manual_seed(77)                                    # same dropout seed for both runs
        res[bits] = _util.run_fwd_bwd(model, bmg, d_out)
        assert bool(enc.last_config.flags & _capi.FLAG_NO_SIGN_BITS) == (bits == "0")
    assert torch.equal(res["1"][0], res["0"][0])
    assert len(res["1"][1]) >= 3
    for k in res["1"][1]:
        assert torch.equal(res["1"][1][k], res["0"][1][k]), k


@pytest.mark.parametrize("mode", ["tf32x3", "tf32", "bf16x3"])
def test_cta_pair_kernel_gives_the_same_bits_as_the_single_cta_kernel(mode, monkeypatch):
    """WD_T4_PAIR=1 runs the forward contractions as clusters of two CTAs (tcgen05.mma.cta_group::2: one M = 256 MMA
    per pair, each CTA holding half of every weight unit; csrc/tc_gemm4.cuh).  Same products accumulated in the same
    order per output element => bitwise the same output and gradients as the default single-CTA kernel.  (Opt-in: it
    measured slower, profiles/r2_fwd_abl.log section F.)  The batch has >= 2 * 148 row tiles, the kernel's threshold,
    and an odd number of 128-row tiles, so one pair's second CTA has no rows at all."""
    bmg = BatchMolGraph(synth.make_graphs("copolymer", 700, 17, random_edge_weights=True))
    n_tiles = (bmg.n_bonds + 127) // 128
    assert n_tiles >= 296
    torch.manual_seed(4)
    model = _util.MPN(_util.make_args(hidden_size=300, depth=3, wd_math_mode=mode), atom_fdim=bmg.atom_fdim,
                      bond_fdim=bmg.bond_fdim).to(DEV)
    d_out = torch.randn(len(bmg.a_scope), 300, generator=torch.Generator().manual_seed(5))
    res = {}
    for flag in ("0", "1"):
        monkeypatch.setenv("WD_T4_PAIR", flag)
        monkeypatch.setenv("WD_GRAPH", "0")
        res[flag] = _util.run_fwd_bwd(model, bmg, d_out)
    (o0, g0), (o1, g1) = res["0"], res["1"]
    assert torch.isfinite(o1).all() and torch.equal(o0, o1)
    for k in g0:
        assert torch.equal(g0[k], g1[k]), k
