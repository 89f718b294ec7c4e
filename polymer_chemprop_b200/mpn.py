"""Drop-in ``MPN`` / ``MPNEncoder`` (chemprop/models/mpn.py:14-289) backed by libwdmpnn.so.

Same constructor arguments, attribute names, sub-module / parameter names
(``encoder.<i>.W_i|W_h|W_o``, ``cached_zero_vector``) and ``forward`` signatures as the
reference, so ``MoleculeModel`` (chemprop/models/model.py:41-47) and existing checkpoints
(chemprop/utils.py:98-135, :194) work unchanged.  The arithmetic runs in hand-written sm_100a
kernels through the C ABI; there is no CPU / eager fallback.
"""
import ctypes as C
import os
from functools import reduce
from typing import List, Optional

import numpy as np
import torch
import torch.nn as nn

from . import _capi
from .graph import ATOM_FDIM, BOND_ONLY_FDIM, BatchMolGraph


def get_activation_function(activation: str) -> nn.Module:
    """chemprop/nn_utils.py:70-99."""
    table = {"ReLU": nn.ReLU, "LeakyReLU": lambda: nn.LeakyReLU(0.1), "PReLU": nn.PReLU,
             "tanh": nn.Tanh, "SELU": nn.SELU, "ELU": nn.ELU}
    if activation not in table:
        raise ValueError(f'Activation "{activation}" not supported.')
    return table[activation]()


_PINNED_FLAGS = []     # recycled 4-int pinned buffers: a fresh cudaHostAlloc per batch costs ~1 ms of host time


class DeviceGraph:
    """Device-resident form of one BatchMolGraph: the get_components() tensors moved once, plus the
    int32 CSR / reverse-CSR arrays rebuilt on the device by wd_csr_build.  Cached on the
    BatchMolGraph instance (like the reference caches a2a / b2b, featurization.py:812-813)."""

    def __init__(self, mol_graph, device: torch.device, validate: bool = True, stream=None):
        """``stream``: build on that CUDA stream without blocking the host (pinned host tensors are
        copied with non_blocking=True); consumers wait on ``self.ready``.  Validation of the graph
        invariants is deferred to ``check()`` (first use), which only waits for this graph's own work."""
        with _capi.device_guard(device):     # the library launches on the current device (the reference never sets it)
            if stream is not None:
                with torch.cuda.stream(stream):
                    self._build(mol_graph, device, non_blocking=True)
            else:
                self._build(mol_graph, device, non_blocking=False)
        self._validate = validate
        self._checked = False

    # Device buffers of a packed build, in arena order: (name, dtype, shape as a function of the sizes)
    _PACKED_OUT = ("f_atoms", "f_bonds", "in_ptr", "in_idx", "out_idx", "src", "rev", "dst", "mol_of", "flags")

    def _build_packed(self, mol_graph, device) -> bool:
        """The upload of a batch that was packed into one pinned arena (BatchMolGraph.pack_arena): ONE host->device
        copy, ONE device allocation for everything that is built (padded feature matrices, CSR arrays, flags), the
        same two C calls.  Bit-identical to the tensor-by-tensor build (tests/test_gpu_parity.py); at batch 50 it takes
        the host ~50 us instead of ~260 us (profiles/r2_c_b50_e2e_profile.log).  Tensor attributes (``f_atoms``,
        ``in_ptr`` ...) are views created on first access."""
        arena, lay = getattr(mol_graph, "_arena", None), getattr(mol_graph, "_arena_layout", None)
        if arena is None or device.type != "cuda" or not arena.is_pinned():
            return False
        Fa, Fb = mol_graph.f_atoms.shape[1], mol_graph.f_bonds.shape[1]
        Na, Nb, K = mol_graph.f_atoms.shape[0], mol_graph.f_bonds_tail.shape[0], mol_graph.a2b.shape[1]
        G = mol_graph.dop_array.numel()
        if mol_graph.f_bonds_tail.shape[1] != Fb - Fa or mol_graph.a_scope_array.numel() != 2 * G:
            return False
        self.device, self.atom_fdim, self.bond_fdim = device, Fa, Fb
        self.n_atoms, self.n_bonds, self.n_mols = Na, Nb, G
        self.compact_upload = True
        d_in = arena.to(device, non_blocking=True)                       # the one upload
        ld_a, ld_b = (Fa + 3) // 4 * 4, (Fb + 3) // 4 * 4
        shapes = {"f_atoms": (torch.float32, (Na, ld_a)), "f_bonds": (torch.float32, (Nb, ld_b)),
                  "in_ptr": (torch.int32, (Na + 1,)), "in_idx": (torch.int32, (max(Nb - 1, 1),)),
                  "out_idx": (torch.int32, (max(Nb - 1, 1),)), "src": (torch.int32, (Nb,)), "rev": (torch.int32, (Nb,)),
                  "dst": (torch.int32, (Nb,)), "mol_of": (torch.int32, (Na,)), "flags": (torch.int32, (4,))}
        views, off = {}, 0
        for name in self._PACKED_OUT:
            dt, shp = shapes[name]
            n = 4
            for d in shp:
                n *= d
            views[name] = (1, off, n, dt, shp)
            off += (n + 255) // 256 * 256
        d_out = torch.empty(off, dtype=torch.uint8, device=device)
        host_shapes = {"w_atoms": (torch.float32, (Na,)), "w_bonds": (torch.float32, (Nb,)),
                       "a_start": (torch.int32, (G,)), "a_size": (torch.int32, (G,)), "dop": (torch.float32, (G,))}
        so = lay["a_scope_array"][0]
        in_off = {"w_atoms": lay["w_atoms"][0], "w_bonds": lay["w_bonds"][0], "a_start": so, "a_size": so + 4 * G,
                  "dop": lay["dop_array"][0]}
        for name, (dt, shp) in host_shapes.items():
            views[name] = (0, in_off[name], 4 * shp[0], dt, shp)
        self._arenas, self._views = (d_in, d_out), views
        bi, bo = d_in.data_ptr(), d_out.data_ptr()
        P = {name: (bo if v[0] else bi) + v[1] for name, v in views.items()}
        L = _capi.lib()
        stream = _capi.current_stream(device)
        _capi.check(L.wd_expand_features(bi + lay["f_atoms"][0], bi + lay["f_bonds_tail"][0], bi + lay["b2a"][0], Na, Nb, Fa,
                                         Fb - Fa, P["f_atoms"], ld_a, P["f_bonds"], ld_b, stream))
        _capi.check(L.wd_csr_build(bi + lay["a2b"][0], Na, K, bi + lay["b2a"][0], bi + lay["b2revb"][0], Nb, P["a_start"],
                                   P["a_size"], G, P["in_ptr"], P["in_idx"], P["out_idx"], P["src"], P["rev"], P["dst"],
                                   P["mol_of"], P["flags"], stream))
        self._dense_keepalive = self._int64_keepalive = None
        self._host_keepalive = (arena,)                                 # pinned source of the in-flight copy
        self.flags_host = _PINNED_FLAGS.pop() if _PINNED_FLAGS else torch.empty(4, dtype=torch.int32).pin_memory()
        self.flags_host.copy_(self.flags, non_blocking=True)
        self.build_stream = torch.cuda.current_stream(device)
        self.ready = torch.cuda.Event()
        self.ready.record(self.build_stream)
        self.struct = _capi.WdGraph(Na, Nb, G, P["f_atoms"], P["f_bonds"], P["w_atoms"], P["w_bonds"], P["in_ptr"],
                                    P["in_idx"], P["out_idx"], P["src"], P["rev"], P["dst"], P["mol_of"], P["a_start"],
                                    P["a_size"], P["dop"], ld_a, ld_b)
        return True

    def __getattr__(self, name):
        # tensor attributes of a packed build: views of the two arenas, made on first access
        views = self.__dict__.get("_views")
        if views is not None and name in views:
            which, off, n, dt, shp = views[name]
            t = self._arenas[which][off:off + n].view(dt).view(shp)
            self.__dict__[name] = t
            return t
        raise AttributeError(name)

    def _build(self, mol_graph, device, non_blocking):
        if self._build_packed(mol_graph, device):
            return
        f_atoms, f_bonds, w_atoms, w_bonds, a2b, b2a, b2revb, a_scope, _b_scope, dop = \
            mol_graph.get_components(atom_messages=False)
        f32 = dict(device=device, dtype=torch.float32, non_blocking=non_blocking)
        i64 = dict(device=device, dtype=torch.int64, non_blocking=non_blocking)
        self.device = device
        tail = getattr(mol_graph, "f_bonds_tail", None)
        self.atom_fdim, self.bond_fdim = f_atoms.shape[1], f_bonds.shape[1]
        self.w_atoms = w_atoms.to(**f32).contiguous()
        self.w_bonds = w_bonds.to(**f32).contiguous()
        a2b = a2b.to(**i64).contiguous()
        b2a = b2a.to(**i64).contiguous()
        b2revb = b2revb.to(**i64).contiguous()
        self.compact_upload = (tail is not None and f_bonds.device.type == "cpu" and
                               tail.shape == (f_bonds.shape[0], self.bond_fdim - self.atom_fdim))
        if self.compact_upload:
            # f_bonds[b] = [f_atoms[b2a[b]], tail[b]] (featurization.py:467-468): upload f_atoms + the bond-only
            # tail and rebuild both padded matrices in HBM (wd_expand_features) -- 65 MB instead of 209 MB at B=4096
            fa_dense = f_atoms.to(**f32).contiguous()
            tail_d = tail.to(**f32).contiguous()
            ld_a, ld_b = (self.atom_fdim + 3) // 4 * 4, (self.bond_fdim + 3) // 4 * 4
            self.f_atoms = torch.empty((fa_dense.shape[0], ld_a), dtype=torch.float32, device=device)
            self.f_bonds = torch.empty((tail_d.shape[0], ld_b), dtype=torch.float32, device=device)
            _capi.check(_capi.lib().wd_expand_features(
                _capi.ptr(fa_dense), _capi.ptr(tail_d), _capi.ptr(b2a), fa_dense.shape[0], tail_d.shape[0],
                self.atom_fdim, self.bond_fdim - self.atom_fdim, _capi.ptr(self.f_atoms), ld_a,
                _capi.ptr(self.f_bonds), ld_b, _capi.current_stream(device)))
            self._dense_keepalive = (fa_dense, tail_d)             # read by the kernels still in flight
        else:
            self.f_atoms = self._padded_rows(f_atoms.to(**f32))    # mpn.py:84-87
            self.f_bonds = self._padded_rows(f_bonds.to(**f32))
            self._dense_keepalive = None
        self.n_atoms, self.n_bonds, self.n_mols = self.f_atoms.shape[0], self.f_bonds.shape[0], len(a_scope)
        if a2b.dim() != 2 or a2b.shape[0] != self.n_atoms:
            raise ValueError("a2b must be [n_atoms, max_num_bonds]")
        host, dop_h = getattr(mol_graph, "a_scope_array", None), getattr(mol_graph, "dop_array", None)
        if host is None or dop_h is None or host.numel() != 2 * len(a_scope) or dop_h.numel() != len(a_scope):
            # a foreign batch object: the reference's Python lists (featurization.py:795-800) are converted here
            scope = np.fromiter((v for se in a_scope for v in se), dtype=np.int64, count=2 * len(a_scope)).reshape(-1, 2)
            host = torch.from_numpy(np.concatenate([scope[:, 0], scope[:, 1]]).astype(np.int32))
            dop_h = torch.from_numpy(np.fromiter((float(x) for x in dop), dtype=np.float32, count=len(dop)))
        host, dop_h = host.to(torch.int32), dop_h.to(torch.float32)              # no-ops for the cached arrays
        if device.type == "cuda":
            host = host if host.is_pinned() else host.pin_memory()
            dop_h = dop_h if dop_h.is_pinned() else dop_h.pin_memory()
        self._host_keepalive = (host, dop_h)                   # pinned sources of in-flight async copies
        scopes = host.to(device, non_blocking=True)
        self.a_start, self.a_size = scopes[:self.n_mols], scopes[self.n_mols:]
        self.dop = dop_h.to(device, non_blocking=True)
        i32 = dict(device=device, dtype=torch.int32)
        self.in_ptr = torch.empty(self.n_atoms + 1, **i32)
        self.in_idx = torch.empty(max(self.n_bonds - 1, 1), **i32)
        self.out_idx = torch.empty(max(self.n_bonds - 1, 1), **i32)
        self.src = torch.empty(self.n_bonds, **i32)
        self.rev = torch.empty(self.n_bonds, **i32)
        self.dst = torch.empty(self.n_bonds, **i32)
        self.mol_of = torch.empty(self.n_atoms, **i32)
        self.flags = torch.empty(4, **i32)
        L, p = _capi.lib(), _capi.ptr
        stream = _capi.current_stream(device)
        _capi.check(L.wd_csr_build(p(a2b), self.n_atoms, a2b.shape[1], p(b2a), p(b2revb), self.n_bonds,
                                   p(self.a_start), p(self.a_size), self.n_mols, p(self.in_ptr),
                                   p(self.in_idx), p(self.out_idx), p(self.src), p(self.rev), p(self.dst),
                                   p(self.mol_of), p(self.flags), stream))
        self._int64_keepalive = (a2b, b2a, b2revb)             # read by the CSR kernels still in flight
        self.flags_host = None
        if device.type == "cuda":
            self.flags_host = _PINNED_FLAGS.pop() if _PINNED_FLAGS else torch.empty(4, dtype=torch.int32).pin_memory()
            self.flags_host.copy_(self.flags, non_blocking=True)
        self.build_stream = torch.cuda.current_stream(device)
        self.ready = torch.cuda.Event()
        self.ready.record(self.build_stream)
        self.struct = _capi.WdGraph(
            self.n_atoms, self.n_bonds, self.n_mols, p(self.f_atoms), p(self.f_bonds), p(self.w_atoms),
            p(self.w_bonds), p(self.in_ptr), p(self.in_idx), p(self.out_idx), p(self.src), p(self.rev),
            p(self.dst), p(self.mol_of), p(self.a_start), p(self.a_size), p(self.dop),
            self.f_atoms.shape[1], self.f_bonds.shape[1])

    @staticmethod
    def _padded_rows(x: torch.Tensor) -> torch.Tensor:
        """Device copy of a feature matrix whose row stride is a multiple of 4 floats (WdGraph.ld_atoms /
        ld_bonds): 133- and 147-float rows are only 4-byte aligned, padded rows take the 16-byte
        asynchronous-copy path of the tensor-core kernels.  Pad columns are zero and never read as data."""
        f = x.shape[1]
        ld = (f + 3) // 4 * 4
        if ld == f:
            return x.contiguous()
        out = torch.zeros((x.shape[0], ld), dtype=x.dtype, device=x.device)
        out[:, :f].copy_(x)
        return out

    def tensors(self):
        if self.__dict__.get("_arenas") is not None:
            return self._arenas
        return (self.f_atoms, self.f_bonds, self.w_atoms, self.w_bonds, self.a_start, self.dop, self.in_ptr,
                self.in_idx, self.out_idx, self.src, self.rev, self.dst, self.mol_of, self.flags)

    def use_on(self, stream):
        """Orders ``stream`` after this graph's build and tells the caching allocator that ``stream`` reads
        the graph's buffers (they were allocated on the build stream)."""
        stream.wait_event(self.ready)
        if stream != self.build_stream:
            for t in self.tensors():
                t.record_stream(stream)

    def check(self):
        """Waits for this graph's upload + CSR build (not for the consumer stream) and raises
        ValueError if the BatchMolGraph invariants were violated (wd_csr_build flags[0])."""
        if self._checked:
            return
        self.ready.synchronize()
        self._int64_keepalive = None
        self._host_keepalive = None
        self._dense_keepalive = None
        flags = self.flags_host.tolist()
        _PINNED_FLAGS.append(self.flags_host)                  # its copy has completed: hand the buffer on
        self.flags_host = None
        self.violations, self.max_in_degree = flags[0], flags[1]
        self.rev_adjacent = flags[2] == 0                      # reverse bonds are adjacent pairs (featurization.py:470-480)
        self.struct.rev_adjacent = int(self.rev_adjacent)
        if self._validate and flags[0] != 0:
            raise ValueError(f"inconsistent BatchMolGraph: {flags[0]} invariant violations "
                             "(bond id out of range, bond listed in 0 or >1 a2b rows, b2revb not an involution, "
                             "or a_scope out of range)")
        self._checked = True

    @staticmethod
    def prefetch(mol_graph, device: torch.device, stream) -> "DeviceGraph":
        """Starts the host->device upload and the device CSR build of ``mol_graph`` on ``stream`` and
        returns immediately; a later ``MPN.forward([mol_graph])`` picks the cached result up and makes
        its own stream wait for it.  This is how a data loader overlaps the upload of batch i+1 with
        the encoder step of batch i."""
        return DeviceGraph.get(mol_graph, device, stream=stream)

    @staticmethod
    def get(mol_graph, device: torch.device, stream=None) -> "DeviceGraph":
        cache = getattr(mol_graph, "_wd_cache", None)
        if cache is None:
            cache = {}
            try:
                mol_graph._wd_cache = cache
            except AttributeError:
                pass
        key = (device.type, device.index)
        if key not in cache:
            cache[key] = DeviceGraph(mol_graph, device, stream=stream)
        return cache[key]


class _EncoderFunction(torch.autograd.Function):
    """forward = wd_encoder_forward, backward = wd_encoder_backward (deterministic)."""

    @staticmethod
    def forward(ctx, enc, dg, cfg, W_i, b_i, W_h, b_h, W_o, b_o):
        dev = dg.device
        with _capi.device_guard(dev):        # the library launches on the CURRENT device: the reference never calls set_device
            ws_bytes = _capi.lib().wd_workspace_bytes(C.byref(cfg), dg.n_atoms, dg.n_bonds, dg.n_mols)
            if ws_bytes == 0:
                _capi.check(_capi.WD_ERR_INVALID_ARG)
            ws = _capi.empty_high_water("ws", (ws_bytes,), torch.uint8, dev)       # capacity: largest bucket so far (_capi)
            out = torch.empty((dg.n_mols, enc.hidden_size), dtype=torch.float32, device=dev)
            ws_c = [t.detach().contiguous() if t is not None else None for t in (W_i, b_i, W_h, b_h, W_o, b_o)]
            weights = _capi.WdWeights(*[_capi.ptr(t) for t in ws_c])
            _capi.check(_capi.lib().wd_encoder_forward(C.byref(dg.struct), C.byref(weights), C.byref(cfg),
                                                       _capi.ptr(ws), ws_bytes, _capi.ptr(out),
                                                       _capi.current_stream(dev)))
        # the parameters themselves are saved (not detached aliases): autograd's version counters then catch an
        # in-place update (an optimizer step) between this forward and its backward
        ctx.present = [t is not None for t in (W_i, b_i, W_h, b_h, W_o, b_o)]
        ctx.save_for_backward(*[t for t in (W_i, b_i, W_h, b_h, W_o, b_o) if t is not None])
        ctx.enc, ctx.dg, ctx.cfg, ctx.ws, ctx.ws_bytes = enc, dg, cfg, ws, ws_bytes
        ctx.has_bias = b_i is not None
        return out

    @staticmethod
    def backward(ctx, d_out):
        enc, dg, cfg = ctx.enc, ctx.dg, ctx.cfg
        dev = dg.device
        if ctx.ws is None:
            raise RuntimeError("the B200 encoder's backward ran twice on one forward: the workspace holding the saved "
                               "activations is released after the first backward (retain_graph is not supported)")
        saved = iter(ctx.saved_tensors)                     # raises if a parameter was modified in place since forward
        W_i, b_i, W_h, b_h, W_o, b_o = [next(saved).detach().contiguous() if here else None for here in ctx.present]
        sink = enc._grad_sink
        with _capi.device_guard(dev):
            d_out = d_out.contiguous().float()
            direct = sink is not None and sink.take(enc)    # first backward of this encoder in the step: write in place
            if direct:
                gW_i, gb_i, gW_h, gb_h, gW_o, gb_o = sink.views_for(enc)
            else:
                gW_i, gW_h, gW_o, gb_o = (torch.empty_like(t) for t in (W_i, W_h, W_o, b_o))
                gb_i = torch.empty_like(b_i) if ctx.has_bias else None
                gb_h = torch.empty_like(b_h) if ctx.has_bias else None
            weights = _capi.WdWeights(*[_capi.ptr(t) for t in (W_i, b_i, W_h, b_h, W_o, b_o)])
            grads = _capi.WdGrads(*[_capi.ptr(t) for t in (gW_i, gb_i, gW_h, gb_h, gW_o, gb_o)])
            _capi.check(_capi.lib().wd_encoder_backward(C.byref(dg.struct), C.byref(weights), C.byref(cfg),
                                                        _capi.ptr(ctx.ws), ctx.ws_bytes, _capi.ptr(d_out),
                                                        C.byref(grads), _capi.current_stream(dev)))
        ctx.ws = None
        if enc.depth == 1:          # W_h is unused by the reference's graph when depth == 1
            if direct:
                gW_h.zero_()
                if gb_h is not None:
                    gb_h.zero_()
            gW_h, gb_h = None, None
        if direct:                  # the gradients already sit in the flat buffer and are the parameters' .grad
            return (None,) * 9
        return None, None, None, gW_i, gb_i, gW_h, gb_h, gW_o, gb_o


class MPNEncoder(nn.Module):
    """Drop-in for chemprop.models.mpn.MPNEncoder (mpn.py:14-173)."""

    def __init__(self, args, atom_fdim: int, bond_fdim: int):
        super().__init__()
        self.atom_fdim = atom_fdim
        self.bond_fdim = bond_fdim
        self.atom_messages = args.atom_messages
        self.hidden_size = args.hidden_size
        self.bias = args.bias
        self.depth = args.depth
        self.dropout = args.dropout
        self.layers_per_message = 1
        self.undirected = args.undirected
        self.device = args.device
        self.aggregation = args.aggregation
        self.aggregation_norm = args.aggregation_norm
        self.math_mode = getattr(args, "wd_math_mode", None) or _default_math_mode()
        self._grad_sink = None      # dp.FlatGradAllReduce.attach(): the backward writes into the flat gradient buffer
        if self.atom_messages:
            raise NotImplementedError("atom_messages=True is not accelerated by the B200 encoder; "
                                      "use chemprop.models.mpn.MPNEncoder for that mode")
        if getattr(args, "atom_descriptors", None) == "descriptor":
            raise NotImplementedError("atom_descriptors='descriptor' is not accelerated by the B200 encoder; "
                                      "use chemprop.models.mpn.MPNEncoder for that mode")
        if args.activation not in _capi.ACTIVATIONS:
            raise NotImplementedError(f'activation "{args.activation}" is not supported by the B200 encoder '
                                      f"(supported: {sorted(_capi.ACTIVATIONS)})")
        if self.aggregation not in _capi.AGGREGATIONS:
            raise ValueError(f'aggregation "{self.aggregation}" not supported')
        if self.math_mode not in _capi.MATH_MODES:
            raise ValueError(f'wd_math_mode "{self.math_mode}" not in {sorted(_capi.MATH_MODES)}')
        self.activation = args.activation
        self.dropout_layer = nn.Dropout(p=self.dropout)
        self.act_func = get_activation_function(args.activation)
        self.cached_zero_vector = nn.Parameter(torch.zeros(self.hidden_size), requires_grad=False)
        self.W_i = nn.Linear(self.bond_fdim, self.hidden_size, bias=self.bias)
        self.W_h = nn.Linear(self.hidden_size, self.hidden_size, bias=self.bias)
        self.W_o = nn.Linear(self.atom_fdim + self.hidden_size, self.hidden_size)

    def _config(self, dg: DeviceGraph) -> _capi.WdConfig:
        """WdConfig of THIS call.  nn.Dropout semantics (mpn.py:38, :124, :134): active only in training mode;
        a fresh 64-bit Philox key is drawn from torch's CPU generator per call (so torch.manual_seed makes runs
        repeatable) and the same cfg object goes to the backward call, which regenerates the masks."""
        p, lo, hi = 0.0, 0, 0
        if self.training and self.dropout > 0.0:
            p = float(self.dropout)
            lo, hi = (int(x) & 0xFFFFFFFF for x in torch.randint(0, 2 ** 32, (2,), dtype=torch.int64).tolist())
        flags = 0
        if os.environ.get("WD_SIGN_BITS", "1") == "0":      # A/B switch; part of the config: forward and backward agree
            flags |= _capi.FLAG_NO_SIGN_BITS
        if os.environ.get("WD_GRAPH", "1") == "0":
            flags |= _capi.FLAG_NO_GRAPH
        if os.environ.get("WD_SIDE_BRANCH", "1") == "0":
            flags |= _capi.FLAG_NO_SIDE_BRANCH
        if os.environ.get("WD_DEFER_SUM", "1") == "0":
            flags |= _capi.FLAG_NO_DEFER_SUM
        return _capi.WdConfig(self.hidden_size, dg.atom_fdim, dg.bond_fdim, self.depth,
                              _capi.ACTIVATIONS[self.activation], _capi.AGGREGATIONS[self.aggregation],
                              int(bool(self.undirected)), int(bool(self.bias)),
                              _capi.MATH_MODES[self.math_mode], float(self.aggregation_norm), p, lo, hi, flags)

    def _infer(self, dg: DeviceGraph, cfg: _capi.WdConfig) -> torch.Tensor:
        """Forward only (torch.no_grad(): chemprop/train/predict.py:10-68, molecule_fingerprint.py:151-183):
        wd_encoder_infer keeps nothing for a backward pass, so the workspace is a third of the training one."""
        L, dev = _capi.lib(), dg.device
        with _capi.device_guard(dev):
            ws_bytes = L.wd_infer_workspace_bytes(C.byref(cfg), dg.n_atoms, dg.n_bonds, dg.n_mols)
            if ws_bytes == 0:
                _capi.check(_capi.WD_ERR_INVALID_ARG)
            ws = _capi.empty_high_water("ws_infer", (ws_bytes,), torch.uint8, dev)  # capacity: largest bucket so far (_capi)
            out = torch.empty((dg.n_mols, self.hidden_size), dtype=torch.float32, device=dev)
            ws_c = [t.detach().contiguous() if t is not None else None
                    for t in (self.W_i.weight, self.W_i.bias, self.W_h.weight, self.W_h.bias, self.W_o.weight, self.W_o.bias)]
            weights = _capi.WdWeights(*[_capi.ptr(t) for t in ws_c])
            _capi.check(L.wd_encoder_infer(C.byref(dg.struct), C.byref(weights), C.byref(cfg), _capi.ptr(ws), ws_bytes,
                                           _capi.ptr(out), _capi.current_stream(dev)))
        return out

    def forward(self, mol_graph, atom_descriptors_batch: List[np.ndarray] = None) -> torch.Tensor:
        """Encodes a batch of molecular graphs -> [num_molecules, hidden_size] (mpn.py:66-173)."""
        if atom_descriptors_batch is not None:
            raise NotImplementedError("atom descriptors are not accelerated by the B200 encoder")
        device = self.W_i.weight.device
        if device.type != "cuda":
            raise RuntimeError("the B200 encoder runs on CUDA devices only (no CPU fallback); "
                               f"parameters are on {device}")
        dg = DeviceGraph.get(mol_graph, device)
        dg.use_on(torch.cuda.current_stream(device))
        dg.check()
        if dg.atom_fdim != self.atom_fdim or dg.bond_fdim != self.bond_fdim:
            raise ValueError(f"feature dims of the batch ({dg.atom_fdim}, {dg.bond_fdim}) do not match the "
                             f"encoder ({self.atom_fdim}, {self.bond_fdim})")
        cfg = self._config(dg)
        self.last_config = cfg                  # introspection (tests read the dropout key of the call)
        params = (self.W_i.weight, self.W_i.bias, self.W_h.weight, self.W_h.bias, self.W_o.weight, self.W_o.bias)
        if not torch.is_grad_enabled() or not any(p is not None and p.requires_grad for p in params):
            return self._infer(dg, cfg)
        return _EncoderFunction.apply(self, dg, cfg, *params)


def _default_math_mode() -> str:
    """bf16x3: tcgen05 tensor cores with error-compensated forward contractions, operands split at bfloat16 and three
    kind::f16 products per element (a_hi w_hi + a_lo w_hi + a_hi w_lo, fp32 accumulation in TMEM) -- encoder output
    4e-6 .. 6e-6 at h=300, predictions <= 2e-5 (the bound BASELINE.json states for a tensor-core path is 1e-4), and the
    fastest of the modes (fewer bytes through shared memory than single-pass TF32 needs).
    WDMPNN_MATH=tf32x3 (split at TF32: 2e-6), tf32 (single pass, ~5e-4 on predictions) or fp32 (CUDA cores) select the
    other modes."""
    return os.environ.get("WDMPNN_MATH", "bf16x3")


class MPN(nn.Module):
    """Drop-in for chemprop.models.mpn.MPN (mpn.py:176-289)."""

    def __init__(self, args, atom_fdim: int = None, bond_fdim: int = None):
        super().__init__()
        if (atom_fdim is None or bond_fdim is None) and (args.overwrite_default_atom_features or
                                                         args.overwrite_default_bond_features):
            raise NotImplementedError("overwrite_default_*_features needs explicit atom_fdim / bond_fdim "
                                      "(the reference resolves them from featurisation globals)")
        self.atom_fdim = atom_fdim or ATOM_FDIM                      # get_atom_fdim(), featurization.py
        self.bond_fdim = bond_fdim or (self.atom_fdim + BOND_ONLY_FDIM)
        self.features_only = args.features_only
        self.use_input_features = args.use_input_features
        self.device = args.device
        self.atom_descriptors = args.atom_descriptors
        self.overwrite_default_atom_features = args.overwrite_default_atom_features
        self.overwrite_default_bond_features = args.overwrite_default_bond_features
        if self.features_only:
            return
        if args.mpn_shared:                                         # mpn.py:204-208
            self.encoder = nn.ModuleList([MPNEncoder(args, self.atom_fdim, self.bond_fdim)] *
                                         args.number_of_molecules)
        else:
            self.encoder = nn.ModuleList([MPNEncoder(args, self.atom_fdim, self.bond_fdim)
                                          for _ in range(args.number_of_molecules)])

    def forward(self, batch, features_batch: List[np.ndarray] = None,
                atom_descriptors_batch: List[np.ndarray] = None,
                atom_features_batch: List[np.ndarray] = None,
                bond_features_batch: List[np.ndarray] = None) -> torch.Tensor:
        if not hasattr(batch[0], "get_components"):                 # mpn.py:229-264 (SMILES / RDKit mols)
            try:
                from chemprop.features import mol2graph             # needs the reference + RDKit
            except Exception as e:
                raise NotImplementedError("featurising SMILES / RDKit molecules needs chemprop.features."
                                          "mol2graph (RDKit); pass BatchMolGraph objects instead") from e
            batch = [[mols[i] for mols in batch] for i in range(len(batch[0]))]
            if self.atom_descriptors == "feature" or bond_features_batch is not None:
                if len(batch) > 1:
                    raise NotImplementedError("Atom/bond descriptors are currently only supported with one "
                                              "molecule per input (i.e., number_of_molecules = 1).")
                batch = [mol2graph(mols=b, atom_features_batch=atom_features_batch,
                                   bond_features_batch=bond_features_batch,
                                   overwrite_default_atom_features=self.overwrite_default_atom_features,
                                   overwrite_default_bond_features=self.overwrite_default_bond_features)
                         for b in batch]
            else:
                batch = [mol2graph(b) for b in batch]
        if self.use_input_features:                                 # mpn.py:266-270
            features_batch = torch.from_numpy(np.stack(features_batch)).float().to(self.device)
            if self.features_only:
                return features_batch
        if self.atom_descriptors == "descriptor":
            raise NotImplementedError("atom_descriptors='descriptor' is not accelerated by the B200 encoder")
        encodings = [enc(ba) for enc, ba in zip(self.encoder, batch)]          # :279
        output = reduce(lambda x, y: torch.cat((x, y), dim=1), encodings)      # :281
        if self.use_input_features:                                 # :283-287
            if len(features_batch.shape) == 1:
                features_batch = features_batch.view(1, -1)
            output = torch.cat([output, features_batch.to(output.device)], dim=1)
        return output


def install(batch_graph: bool = True):
    """Monkey-patches an importable chemprop so MoleculeModel builds this MPN
    (chemprop/models/model.py:8 did ``from .mpn import MPN``).

    batch_graph: also put this package's BatchMolGraph where the reference's loader builds its batches
    (``MoleculeDataset.batch_graph``, chemprop/data/data.py:312, and ``mol2graph``, features/featurization.py:878-898).
    It accepts the reference's MolGraph objects, fills the same attributes and the same get_components() 10-tuple,
    and additionally offers the compact upload form (``f_bonds_tail``, array scopes): 77 MB instead of 220 MB cross
    PCIe per 4096-graph batch, and the batch is assembled with vectorised numpy instead of Python list appends."""
    import chemprop.models.model as cm
    import chemprop.models.mpn as cmpn
    cmpn.MPN, cmpn.MPNEncoder = MPN, MPNEncoder
    cm.MPN = MPN
    if batch_graph:
        import importlib
        for name in ("chemprop.features.featurization", "chemprop.features", "chemprop.data.data", "chemprop.models.model",
                     "chemprop.models.mpn"):
            try:
                mod = importlib.import_module(name)
            except Exception:          # e.g. chemprop.data needs RDKit: nothing to patch there then
                continue
            if hasattr(mod, "BatchMolGraph"):
                mod.BatchMolGraph = BatchMolGraph
