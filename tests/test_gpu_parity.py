"""GPU parity tests proper: the CUDA path (through the C ABI) against the reference's golden
vectors and against the CPU oracle on seeded synthetic inputs.

Tolerances (BASELINE.json north_star): indices bit-exact; FP32 path <= 1e-5 relative on the
encoder output (||x-ref||_F / ||ref||_F), <= 1e-4 on gradients (summation order differs)."""
import ctypes as C

import numpy as np
import pytest
import torch

import _golden
import _util
from oracle import mpn_oracle
from polymer_chemprop_b200 import MPN, BatchMolGraph, DeviceGraph, _capi, synth

pytestmark = pytest.mark.gpu
DEV = torch.device("cuda", 0)
OUT_TOL, GRAD_TOL = 1e-5, 1e-4


def _check_csr(components):
    g = _util.ComponentsGraph(components)
    dg = DeviceGraph(g, DEV)
    a2b, b2a, b2revb = components[4], components[5], components[6]
    in_ptr, in_idx, out_ptr, out_idx, dst = mpn_oracle.csr_from_a2b(a2b.tolist(), b2a.tolist(), b2revb.tolist())
    nb1 = b2a.numel() - 1
    assert np.array_equal(dg.in_ptr.cpu().numpy(), in_ptr)
    assert np.array_equal(dg.in_ptr.cpu().numpy(), out_ptr)          # out-degree == in-degree
    assert np.array_equal(dg.in_idx.cpu().numpy()[:nb1], in_idx)
    assert np.array_equal(dg.out_idx.cpu().numpy()[:nb1], out_idx)
    assert np.array_equal(dg.dst.cpu().numpy(), dst)
    assert np.array_equal(dg.src.cpu().numpy(), b2a.numpy().astype(np.int32))
    assert np.array_equal(dg.rev.cpu().numpy(), b2revb.numpy().astype(np.int32))
    mol_of = np.full(a2b.shape[0], -1, np.int32)
    for i, (s, n) in enumerate(components[7]):
        mol_of[s:s + n] = i
    assert np.array_equal(dg.mol_of.cpu().numpy(), mol_of)
    flags = dg.flags.cpu().numpy()
    assert flags[0] == 0 and flags[1] == max(1, int((a2b != 0).sum(1).max())) or flags[1] == int((a2b != 0).sum(1).max())


@pytest.mark.parametrize("name", _golden.NAMES)
def test_csr_bit_exact_golden(name):
    _check_csr(_golden.load(name)["components"])


def test_csr_bit_exact_large():
    bmg = BatchMolGraph(synth.make_graphs("copolymer", 700, 5) + [synth.empty_graph()] +
                        synth.make_graphs("qm9", 300, 6))
    _check_csr(bmg.get_components())


def test_csr_rejects_inconsistent_graph():
    comps = list(_golden.load("copoly_b2")["components"])
    b2revb = comps[6].clone()
    b2revb[3], b2revb[4] = b2revb[5].item(), b2revb[6].item()       # break the involution
    comps[6] = b2revb
    with pytest.raises(ValueError):
        DeviceGraph(_util.ComponentsGraph(tuple(comps)), DEV).check()


def test_compact_upload_rebuilds_features_bit_exactly():
    """wd_expand_features: f_atoms + the bond-only tail uploaded, f_bonds rebuilt in HBM == the full upload."""
    graphs = synth.make_graphs("copolymer", 300, 11, dense_features=True) + [synth.empty_graph()] + \
        synth.make_graphs("qm9", 100, 12)
    bmg = BatchMolGraph(graphs)
    assert bmg.f_bonds_tail is not None
    compact = DeviceGraph(bmg, DEV)
    full = DeviceGraph(_util.ComponentsGraph(bmg.get_components()), DEV)
    assert compact.compact_upload and not full.compact_upload
    for name in ("f_atoms", "f_bonds", "w_atoms", "w_bonds", "in_ptr", "in_idx", "out_idx", "src", "rev", "dst", "mol_of"):
        a, b = getattr(compact, name), getattr(full, name)
        assert a.shape == b.shape and torch.equal(a.view(torch.int32), b.view(torch.int32)), name
    assert torch.equal(compact.f_bonds[:, :bmg.bond_fdim].cpu(), bmg.f_bonds)
    assert torch.equal(compact.f_atoms[:, :bmg.atom_fdim].cpu(), bmg.f_atoms)
    # a stream-ordered build (the data-loader path) gives the same bytes
    side = torch.cuda.Stream(device=DEV)
    hg = _util.ComponentsGraph(tuple(x.pin_memory() if torch.is_tensor(x) else x for x in bmg.get_components()))
    hg.f_bonds_tail = bmg.f_bonds_tail.pin_memory()
    pre = DeviceGraph.prefetch(hg, DEV, side)
    pre.use_on(torch.cuda.current_stream(DEV))
    pre.check()
    assert pre.compact_upload and torch.equal(pre.f_bonds, full.f_bonds) and torch.equal(pre.f_atoms, full.f_atoms)


def test_packed_upload_is_bit_identical_to_the_tensor_by_tensor_build():
    """BatchMolGraph.pin_memory packs the upload set into ONE pinned arena; DeviceGraph then uploads with a single copy
    and builds everything into a single device allocation.  Same bytes in every array, same encoder output and gradients
    (bitwise), on the current stream and on a copy stream."""
    graphs = synth.make_graphs("copolymer", 120, 5, dense_features=True, random_edge_weights=True) + [synth.empty_graph()] + \
        synth.make_graphs("qm9", 40, 6)
    plain = BatchMolGraph(graphs)
    packed = BatchMolGraph(graphs).pin_memory()
    assert packed._arena is not None and packed._arena.is_pinned() and getattr(plain, "_arena", None) is None
    for name in BatchMolGraph._UPLOAD:
        assert torch.equal(getattr(packed, name), getattr(plain, name)), name
    ref = DeviceGraph(plain, DEV)
    ref.check()
    side = torch.cuda.Stream(device=DEV)
    for dg in (DeviceGraph(packed.fresh_view(), DEV), DeviceGraph.prefetch(packed.fresh_view(), DEV, side)):
        assert dg.__dict__.get("_arenas") is not None and dg.compact_upload
        dg.use_on(torch.cuda.current_stream(DEV))
        dg.check()
        assert (dg.violations, dg.max_in_degree, dg.rev_adjacent) == (ref.violations, ref.max_in_degree, ref.rev_adjacent)
        for name in ("f_atoms", "f_bonds", "w_atoms", "w_bonds", "in_ptr", "in_idx", "out_idx", "src", "rev", "dst", "mol_of",
                     "a_start", "a_size", "dop"):
            a, b = getattr(dg, name), getattr(ref, name)
            assert a.shape == b.shape and a.dtype == b.dtype and torch.equal(a.view(torch.int32), b.view(torch.int32)), name
    torch.manual_seed(2)
    model = MPN(_util.make_args(hidden_size=64, depth=3, wd_math_mode="tf32x3"), atom_fdim=plain.atom_fdim,
                bond_fdim=plain.bond_fdim).to(DEV)
    d_out = torch.randn(len(graphs), 64)
    out_a, grads_a = _util.run_fwd_bwd(model, plain, d_out)
    out_b, grads_b = _util.run_fwd_bwd(model, packed.fresh_view(), d_out)
    assert torch.equal(out_a, out_b) and all(torch.equal(grads_a[k], grads_b[k]) for k in grads_a)


@pytest.mark.parametrize("name", _golden.NAMES)
def test_encoder_matches_reference_golden_fp32(name):
    g = _golden.load(name)
    model = _util.model_from_params(g["params"], g["cfg"])
    out, grads = _util.run_fwd_bwd(model, _util.ComponentsGraph(g["components"]), g["d_out"])
    assert out.shape == g["out"].shape
    assert _golden.rel_err(out, g["out"]) <= OUT_TOL
    assert set(grads) == set(g["grads"])
    for k, ref in g["grads"].items():
        assert _golden.rel_err(grads[k], ref) <= GRAD_TOL, k


@pytest.mark.parametrize("activation", ["ReLU", "LeakyReLU", "tanh", "SELU", "ELU"])
@pytest.mark.parametrize("undirected,bias", [(False, False), (True, True)])
def test_encoder_matches_oracle_synthetic(activation, undirected, bias):
    """Seeded synthetic batch, odd hidden size, every supported activation."""
    graphs = synth.make_graphs("copolymer", 24, 21, dense_features=True, random_edge_weights=True)
    bmg = BatchMolGraph(graphs + [synth.empty_graph()])
    cfg = dict(hidden_size=76, depth=4, activation=activation, undirected=undirected, bias=bias,
               aggregation="mean")
    torch.manual_seed(5)
    model = _util.MPN(_util.make_args(**cfg), atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(DEV)
    for p in model.parameters():
        if p.requires_grad:
            torch.nn.init.normal_(p, 0.0, 0.1) if p.dim() == 1 else torch.nn.init.xavier_normal_(p)
    d_out = torch.randn(len(bmg.a_scope), 76, generator=torch.Generator().manual_seed(9))
    out, grads = _util.run_fwd_bwd(model, bmg, d_out)
    weights = {k: v.detach().cpu() for k, v in model.encoder[0].state_dict().items() if k != "cached_zero_vector"}
    ref_out, ref_grads = mpn_oracle.encoder_forward_backward(
        bmg.get_components(), weights, d_out, depth=4, activation=activation, undirected=undirected)
    assert _golden.rel_err(out, ref_out) <= OUT_TOL
    for k, ref in ref_grads.items():
        assert _golden.rel_err(grads[k], ref) <= GRAD_TOL, k


def test_backward_is_bitwise_deterministic():
    g = _golden.load("copoly_b50")
    model = _util.model_from_params(g["params"], g["cfg"])
    graph = _util.ComponentsGraph(g["components"])
    out1, grads1 = _util.run_fwd_bwd(model, graph, g["d_out"])
    out2, grads2 = _util.run_fwd_bwd(model, graph, g["d_out"])
    assert torch.equal(out1, out2)
    for k in grads1:
        assert torch.equal(grads1[k], grads2[k]), k


def test_host_buffer_entry_point_matches_golden():
    """wd_encoder_host: the reference's own calling convention (CPU tensors in, mpn.py:81-87)."""
    g = _golden.load("copoly_b8_randw")
    fa, fb, wa, wb, a2b, b2a, b2revb, a_scope, _bs, dop = g["components"]
    cfg, P = g["cfg"], g["params"]
    h = cfg["hidden_size"]
    c = _capi.WdConfig(h, fa.shape[1], fb.shape[1], cfg["depth"], 0, 0, 0, 0, 0, 100.0)
    scope = torch.tensor(a_scope, dtype=torch.int64).contiguous()
    dopt = torch.tensor(dop, dtype=torch.float64)
    out = torch.empty(len(a_scope), h)
    grads = {k: torch.empty_like(v) for k, v in P.items() if k != "cached_zero_vector"}
    w = _capi.WdWeights(P["W_i.weight"].data_ptr(), None, P["W_h.weight"].data_ptr(), None,
                        P["W_o.weight"].data_ptr(), P["W_o.bias"].data_ptr())
    gr = _capi.WdGrads(grads["W_i.weight"].data_ptr(), None, grads["W_h.weight"].data_ptr(), None,
                       grads["W_o.weight"].data_ptr(), grads["W_o.bias"].data_ptr())
    torch.cuda.init()
    rc = _capi.lib().wd_encoder_host(fa.data_ptr(), fb.data_ptr(), wa.data_ptr(), wb.data_ptr(), a2b.data_ptr(),
                                     b2a.data_ptr(), b2revb.data_ptr(), scope.data_ptr(), dopt.data_ptr(),
                                     fa.shape[0], fb.shape[0], a2b.shape[1], len(a_scope), C.byref(w),
                                     C.byref(c), out.data_ptr(), g["d_out"].data_ptr(), C.byref(gr), None)
    _capi.check(rc)
    assert _golden.rel_err(out, g["out"]) <= OUT_TOL
    for k, ref in g["grads"].items():
        assert _golden.rel_err(grads[k], ref) <= GRAD_TOL, k


def test_full_size_properties_b4096():
    """BASELINE config 2 at full size (B=4096, h=300, d=3): size-independent properties.
    (1) batch-permutation equivariance of the readout rows; (2) molecules are independent, so the
    first 64 rows equal the encoder run on those 64 graphs alone; (3) gradients of a sum over
    disjoint halves add up (linearity of the backward in d_out)."""
    graphs = synth.make_graphs("copolymer", 4096, 0)
    torch.manual_seed(0)
    bmg = BatchMolGraph(graphs)
    model = _util.MPN(_util.make_args(), atom_fdim=bmg.atom_fdim, bond_fdim=bmg.bond_fdim).to(DEV)
    for p in model.parameters():
        if p.requires_grad and p.dim() == 2:
            torch.nn.init.xavier_normal_(p)
    with torch.no_grad():
        out = model([bmg]).cpu()
        small = model([BatchMolGraph(graphs[:64])]).cpu()
        perm = np.random.default_rng(0).permutation(4096)
        out_p = model([BatchMolGraph([graphs[i] for i in perm])]).cpu()
    assert torch.isfinite(out).all()
    assert _golden.rel_err(out[:64], small) <= 1e-6
    assert _golden.rel_err(out_p, out[perm]) <= 1e-6
    d = torch.randn(4096, 300, generator=torch.Generator().manual_seed(1))
    d1, d2 = d.clone(), d.clone()
    d1[2048:] = 0
    d2[:2048] = 0
    _, g = _util.run_fwd_bwd(model, bmg, d)
    _, g1 = _util.run_fwd_bwd(model, bmg, d1)
    _, g2 = _util.run_fwd_bwd(model, bmg, d2)
    for k in g:
        assert _golden.rel_err(g1[k] + g2[k], g[k]) <= 1e-5, k


@pytest.mark.parametrize("hidden,depth,activation,p", [(76, 4, "ReLU", 0.3), (300, 3, "tanh", 0.15), (44, 2, "ELU", 0.5)])
def test_dropout_matches_oracle_with_the_same_philox_masks_fp32(hidden, depth, activation, p):
    """Training-mode dropout (mpn.py:124, :134): the reference's RNG stream cannot be shared with a CUDA kernel, so
    the oracle applies the product's Philox masks (pinned by Random123's known answers, and its op placement by the
    live reference: tests/test_oracle_vs_reference.py); forward and backward then agree to the FP32 tolerances."""
    _util.dropout_case("fp32", hidden, depth, activation, p, OUT_TOL, GRAD_TOL)


def test_inference_path_equals_training_forward_bitwise():
    """torch.no_grad() -> wd_encoder_infer (predict.py:10-68): same kernels, smaller workspace, same bits."""
    import ctypes as C
    bmg = BatchMolGraph(synth.make_graphs("copolymer", 300, 41) + [synth.empty_graph()])
    for mode in ("fp32", "tf32", "tf32x3", "bf16x3"):
        torch.manual_seed(3)
        model = _util.MPN(_util.make_args(hidden_size=300, depth=4, wd_math_mode=mode), atom_fdim=bmg.atom_fdim,
                          bond_fdim=bmg.bond_fdim).to(DEV)
        out_train = model([bmg])
        assert out_train.requires_grad
        with torch.no_grad():
            out_inf = model([bmg])
        assert not out_inf.requires_grad and torch.equal(out_train.detach(), out_inf)
        for q in model.parameters():                      # frozen encoder (model.py:49-55) takes the same path
            q.requires_grad_(False)
        assert torch.equal(model([bmg]), out_inf)
        cfg = model.encoder[0].last_config
        lib = _capi.lib()
        full = lib.wd_workspace_bytes(C.byref(cfg), bmg.n_atoms, bmg.n_bonds, len(bmg.a_scope))
        small = lib.wd_infer_workspace_bytes(C.byref(cfg), bmg.n_atoms, bmg.n_bonds, len(bmg.a_scope))
        assert 0 < small < 0.45 * full


def _star_graph(rng, n_leaves):
    """One hub atom bonded to n_leaves atoms (in-degree n_leaves: exercises long CSR lists and max_num_bonds >> 8)."""
    n = n_leaves + 1
    fa = synth._atom_features(rng, n, True)
    fb = synth._bond_features(rng, n_leaves, True)
    pairs = [(0, i) for i in range(1, n)]
    w = [(float(rng.uniform(0.05, 1.0)), float(rng.uniform(0.05, 1.0))) for _ in pairs]
    return synth._assemble(n, pairs, w, fa, fb, rng.uniform(0.2, 1.0, n).astype(np.float32), 1.0 + float(rng.uniform(1, 3)))


@pytest.mark.parametrize("mode,out_tol,grad_tol", [("fp32", OUT_TOL, GRAD_TOL), ("tf32x3", 1e-5, 5e-3), ("bf16x3", 2e-5, 5e-3), ("tf32", 1e-3, 3e-2)])
@pytest.mark.parametrize("case", ["depth1", "star_k40", "one_molecule_h8"])
def test_edge_shapes_match_oracle(case, mode, out_tol, grad_tol):
    """Edges of the shape space: depth 1 (no message passing, W_h unused: the reference leaves its grad None),
    an atom with 40 incoming bonds (K = 40), a single molecule with hidden size 8."""
    rng = np.random.default_rng(17)
    if case == "depth1":
        graphs, hidden, depth = synth.make_graphs("copolymer", 6, 81, dense_features=True, random_edge_weights=True), 44, 1
    elif case == "star_k40":
        graphs, hidden, depth = [_star_graph(rng, 40), synth.empty_graph(), _star_graph(rng, 3)], 76, 3
    else:
        graphs, hidden, depth = synth.make_graphs("qm9", 1, 82, dense_features=True), 8, 3
    bmg = BatchMolGraph(graphs)
    if case == "star_k40":
        assert bmg.max_num_bonds == 40
    torch.manual_seed(5)
    model = _util.MPN(_util.make_args(hidden_size=hidden, depth=depth, wd_math_mode=mode), atom_fdim=bmg.atom_fdim,
                      bond_fdim=bmg.bond_fdim).to(DEV)
    for p in model.parameters():
        if p.requires_grad:
            torch.nn.init.normal_(p, 0.0, 0.1) if p.dim() == 1 else torch.nn.init.xavier_normal_(p)
    d_out = torch.randn(len(bmg.a_scope), hidden, generator=torch.Generator().manual_seed(9))
    out, grads = _util.run_fwd_bwd(model, bmg, d_out)
    weights = {k: v.detach().cpu() for k, v in model.encoder[0].state_dict().items() if k != "cached_zero_vector"}
    ref_out, ref_grads = mpn_oracle.encoder_forward_backward(bmg.get_components(), weights, d_out, depth=depth)
    assert _golden.rel_err(out, ref_out) <= out_tol
    for k, ref in ref_grads.items():
        if ref is None:                                   # depth 1: W_h takes no part in the reference's graph
            assert k not in grads
            continue
        assert _golden.rel_err(grads[k], ref) <= grad_tol, k


@pytest.mark.parametrize("mode,depth,bias", [("fp32", 3, False), ("tf32x3", 3, False), ("tf32x3", 5, True), ("tf32x3", 2, True)])
def test_small_batches_run_as_one_cuda_graph_per_pass_with_the_same_bits(mode, depth, bias, monkeypatch):
    """Small batches (the reference's default is 50 graphs, chemprop/args.py:95): the library captures each pass into a
    CUDA graph and re-parameterises ONE executable graph per pass for every new batch (cudaGraphExecUpdate: new
    pointers, sizes, grids and tensor maps).  Same kernels, same order => bitwise the same results as the plain launches
    (WD_GRAPH=0 -> WdConfig.flags & WD_FLAG_NO_GRAPH), for batches of different sizes in a row.  In the captured backward
    pass the weight-gradient kernels sit on a parallel branch of the graph (csrc/encoder.cu: SideBranch; depth 5 exercises
    the join that protects the alternating dZ buffers): still the same bits, also against WD_SIDE_BRANCH=0."""
    batches = [BatchMolGraph(synth.make_graphs("copolymer", n, 200 + n, random_edge_weights=True)) for n in (50, 37, 64)]
    torch.manual_seed(2)
    model = _util.MPN(_util.make_args(hidden_size=300, depth=depth, bias=bias, wd_math_mode=mode),
                      atom_fdim=batches[0].atom_fdim, bond_fdim=batches[0].bond_fdim).to(DEV)
    for p in model.parameters():
        if p.requires_grad and p.dim() == 2:
            torch.nn.init.xavier_normal_(p)
    res = {}
    for flag in ("1", "0", "no_side"):
        monkeypatch.setenv("WD_GRAPH", "0" if flag == "0" else "1")
        monkeypatch.setenv("WD_SIDE_BRANCH", "0" if flag == "no_side" else "1")
        g0, b0 = _capi.graph_stats()
        outs = []
        for bmg in batches:
            d = torch.randn(len(bmg.a_scope), 300, generator=torch.Generator().manual_seed(3))
            outs.append(_util.run_fwd_bwd(model, bmg.fresh_view(), d))
            with torch.no_grad():
                outs.append((model([bmg.fresh_view()]).cpu(), {}))
        g1, b1 = _capi.graph_stats()
        if flag != "0":
            assert g1 - g0 == 9                      # forward, backward, infer for each of the three batches
            assert b1 - b0 <= 3                      # at most one executable graph per pass; the rest were updates
        else:
            assert g1 == g0
        res[flag] = outs
    for other in ("0", "no_side"):
        for (o1, gr1), (o0, gr0) in zip(res["1"], res[other]):
            assert torch.equal(o1, o0)
            for k in gr1:
                assert torch.equal(gr1[k], gr0[k]), (other, k)


@pytest.mark.parametrize("mode", ["fp32", "tf32x3"])
def test_deferred_gradient_sum_gives_the_same_bits_as_the_running_sum(mode, monkeypatch):
    """Depth 3 without a W_h bias: the middle message-gradient step leaves dI = dZ_2 + dZ_1 to the last step, which forms
    (dZ_2 + dZ_1) + dZ_0 from the two dZ buffers (csrc/encoder.cu: defer_sum) -- the same association as the running
    sum, so every gradient agrees bit for bit with WD_DEFER_SUM=0 (WdConfig.flags & WD_FLAG_NO_DEFER_SUM)."""
    bmg = BatchMolGraph(synth.make_graphs("copolymer", 200, 91, random_edge_weights=True) + [synth.empty_graph()])
    torch.manual_seed(6)
    model = MPN(_util.make_args(hidden_size=300, depth=3, wd_math_mode=mode), atom_fdim=bmg.atom_fdim,
                bond_fdim=bmg.bond_fdim).to(DEV)
    d_out = torch.randn(len(bmg.a_scope), 300, generator=torch.Generator().manual_seed(7))
    res = {}
    for flag in ("1", "0"):
        monkeypatch.setenv("WD_DEFER_SUM", flag)
        res[flag] = _util.run_fwd_bwd(model, bmg, d_out)
        assert bool(model.encoder[0].last_config.flags & _capi.FLAG_NO_DEFER_SUM) == (flag == "0")
    assert torch.equal(res["1"][0], res["0"][0])
    for k in res["1"][1]:
        assert torch.equal(res["1"][1][k], res["0"][1][k]), k
