"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference wD-MPNN encoder.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this file; the product package never does.

Parity status: PINNED.  tests/golden/*.npz were generated in the build container by
oracle/make_golden.py from the UNMODIFIED reference module (imported through
oracle/ref_shim.py); tests/test_oracle_golden.py checks this restatement against
them, and tests/test_oracle_vs_reference.py re-runs the live reference when the tree
is present.  (The reference's own tests pin nothing at this boundary)

Two restatements are given:
  * ``encoder_forward`` -- the reference's op sequence on torch CPU tensors (padded
    a2b gather, per-molecule Python readout loop), differentiated by autograd, so that
    timing it reproduces the reference's CPU cost structure (chemprop/models/mpn.py:66-173);
  * ``encoder_forward_backward_explicit`` -- numpy float64, hand-derived backward
    (SURVEY.md section 3.2), used to pin the gradient formulae the CUDA path implements.
  * ``csr_from_a2b`` -- plain-Python index oracle for the device CSR builder.
"""
from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch


def index_select_nd(source: torch.Tensor, index: torch.Tensor) -> torch.Tensor:
    """chemprop/nn_utils.py:50-67: flat index_select then view to index.shape + suffix."""
    flat = source.index_select(0, index.reshape(-1))
    return flat.view(tuple(index.shape) + tuple(source.shape[1:]))


def _act(name: str):
    """chemprop/nn_utils.py:70-99 (stateless activations only; PReLU is rejected by the product)."""
    table = {"ReLU": torch.relu, "LeakyReLU": lambda x: torch.nn.functional.leaky_relu(x, 0.1),
             "tanh": torch.tanh, "SELU": torch.selu, "ELU": torch.nn.functional.elu}
    return table[name]


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Philox4x32-10 (Salmon et al., SC'11) on numpy uint64 arrays holding 32-bit words; the product's
    dropout masks are words of this generator (csrc/common.cuh philox4x32_10)."""
    m32 = np.uint64(0xFFFFFFFF)
    c0, c1, c2, c3 = (np.asarray(x, dtype=np.uint64) & m32 for x in (c0, c1, c2, c3))
    k0, k1 = np.uint64(k0), np.uint64(k1)
    for _ in range(10):
        p0 = np.uint64(0xD2511F53) * c0
        p1 = np.uint64(0xCD9E8D57) * c2
        c0, c1, c2, c3 = ((p1 >> np.uint64(32)) ^ c1 ^ k0) & m32, p1 & m32, ((p0 >> np.uint64(32)) ^ c3 ^ k1) & m32, p0 & m32
        k0 = (k0 + np.uint64(0x9E3779B9)) & m32
        k1 = (k1 + np.uint64(0xBB67AE85)) & m32
    return c0, c1, c2, c3


def dropout_multiplier(n_rows: int, n_cols: int, ld: int, p: float, seed_lo: int, seed_hi: int, salt: int) -> torch.Tensor:
    """The product's dropout mask of one [n_rows, ld]-strided activation as a float32 [n_rows, n_cols] tensor of
    0 or 1/(1-p): element e = row * ld + col is dropped iff word (e & 3) of Philox(counter = (e >> 2, salt, 0),
    key = seed) is below floor(p * 2^32)   (csrc/common.cuh dropout_keep4; nn.Dropout at mpn.py:124 / :134)."""
    e = (np.arange(n_rows, dtype=np.uint64)[:, None] * np.uint64(ld) + np.arange(n_cols, dtype=np.uint64)[None, :])
    q = e >> np.uint64(2)
    words = philox4x32_10(q & np.uint64(0xFFFFFFFF), q >> np.uint64(32), np.uint64(salt), np.uint64(0), seed_lo, seed_hi)
    lane = (e & np.uint64(3)).astype(np.int64)
    w = np.choose(lane, words)
    thresh = np.uint64(min(4294967295, int(np.floor(float(np.float32(p)) * 4294967296.0))))
    scale = np.float32(1.0) / (np.float32(1.0) - np.float32(p))
    return torch.from_numpy(np.where(w < thresh, np.float32(0.0), scale).astype(np.float32))


def encoder_forward(components, weights: Dict[str, torch.Tensor], depth: int,
                    activation: str = "ReLU", undirected: bool = False,
                    aggregation: str = "mean", aggregation_norm: float = 100.0, dropout=None,
                    readout: bool = True) -> torch.Tensor:
    """MPNEncoder.forward, bond-message branch (mpn.py:81-173).  ``dropout`` = None (p = 0 / eval) or a dict
    {p, seed_lo, seed_hi, ld} selecting the product's Philox masks (the reference's own RNG stream cannot be
    shared with a CUDA kernel; the op sequence -- mask * 1/(1-p) after each activation, :124 and :134 -- is the same).

    ``weights`` uses the reference's parameter names: W_i.weight, W_h.weight, W_o.weight,
    W_o.bias and optionally W_i.bias / W_h.bias (mpn.py:47-57).
    """
    f_atoms, f_bonds, w_atoms, w_bonds, a2b, b2a, b2revb, a_scope, _b_scope, dop = components
    act = _act(activation)
    lin = torch.nn.functional.linear
    inp = lin(f_bonds, weights["W_i.weight"], weights.get("W_i.bias"))           # mpn.py:96
    message = act(inp)                                                             # :97
    for _ in range(depth - 1):                                                     # :100  (step t = _ + 1)
        if undirected:
            message = (message + message[b2revb]) / 2                              # :101-102
        nei_msg = index_select_nd(message, a2b)                                    # :112
        nei_w = index_select_nd(w_bonds, a2b)                                      # :113
        a_message = (nei_msg * nei_w[..., None]).sum(dim=1)                        # :117-118
        message = a_message[b2a] - message[b2revb]                                 # :119-120 (reverse unweighted)
        message = lin(message, weights["W_h.weight"], weights.get("W_h.bias"))    # :122
        message = act(inp + message)                                               # :123
        if dropout:                                                                # :124
            message = message * dropout_multiplier(message.shape[0], message.shape[1], dropout["ld"], dropout["p"],
                                                   dropout["seed_lo"], dropout["seed_hi"], _ + 1).to(message.dtype)
    nei_msg = index_select_nd(message, a2b)                                        # :127
    nei_w = index_select_nd(w_bonds, a2b)                                          # :128
    a_message = (nei_msg * nei_w[..., None]).sum(dim=1)                            # :130-131
    a_input = torch.cat([f_atoms, a_message], dim=1)                               # :132
    atom_hiddens = act(lin(a_input, weights["W_o.weight"], weights["W_o.bias"]))  # :133
    if dropout:                                                                    # :134
        atom_hiddens = atom_hiddens * dropout_multiplier(atom_hiddens.shape[0], atom_hiddens.shape[1], dropout["ld"],
                                                         dropout["p"], dropout["seed_lo"], dropout["seed_hi"],
                                                         depth).to(atom_hiddens.dtype)
    if not readout:          # bench.py's "message passing only" CPU figure: stop before mpn.py:146
        return atom_hiddens
    mol_vecs = []
    zero = torch.zeros(atom_hiddens.shape[1], dtype=atom_hiddens.dtype)
    for i, (a_start, a_size) in enumerate(a_scope):                                # :147-169
        if a_size == 0:
            mol_vecs.append(zero)                                                  # cached_zero_vector, unscaled
            continue
        cur = atom_hiddens.narrow(0, a_start, a_size)
        w = w_atoms.narrow(0, a_start, a_size)
        vec = w[..., None] * cur
        if aggregation == "mean":
            vec = vec.sum(dim=0) / w.sum(dim=0)
        elif aggregation == "sum":
            vec = vec.sum(dim=0)
        elif aggregation == "norm":
            vec = vec.sum(dim=0) / aggregation_norm
        mol_vecs.append(dop[i] * vec)
    return torch.stack(mol_vecs, dim=0)


def encoder_forward_backward(components, weights, d_out, **cfg):
    """Forward + autograd backward on CPU; returns (out, grads-by-name)."""
    ws = {k: v.detach().clone().requires_grad_(True) for k, v in weights.items()}
    out = encoder_forward(components, ws, **cfg)
    if not cfg.get("readout", True):
        d_out = torch.ones_like(out)       # backward from atom_hiddens.sum()
    out.backward(d_out)
    return out.detach(), {k: v.grad for k, v in ws.items()}


def csr_from_a2b(a2b: Sequence[Sequence[int]], b2a: Sequence[int], b2revb: Sequence[int]):
    """Index oracle for the device CSR builder, in plain Python.

    in_ptr/in_idx : non-zero entries of each padded a2b row, reference order (SURVEY 3.2:
                    real bond ids are >= 1, so these are exactly the bonds the reference
                    sums with non-zero weight slot, mpn.py:112-118);
    dst           : dst[b] = b2a[b2revb[b]]  (the atom whose a2b row contains b);
    out_ptr/out_idx : bonds grouped by source atom b2a[b] for b >= 1, ascending bond id
                    (transpose of ``a_message[b2a]``, mpn.py:120, used by the backward).
    """
    n_a, n_b = len(a2b), len(b2a)
    in_ptr, in_idx = [0], []
    for row in a2b:
        in_idx.extend(int(b) for b in row if b != 0)
        in_ptr.append(len(in_idx))
    dst = [int(b2a[b2revb[b]]) for b in range(n_b)]
    buckets: List[List[int]] = [[] for _ in range(n_a)]
    for b in range(1, n_b):
        buckets[int(b2a[b])].append(b)
    out_ptr, out_idx = [0], []
    for row in buckets:
        out_idx.extend(row)
        out_ptr.append(len(out_idx))
    i32 = np.int32
    return (np.asarray(in_ptr, i32), np.asarray(in_idx, i32), np.asarray(out_ptr, i32),
            np.asarray(out_idx, i32), np.asarray(dst, i32))
