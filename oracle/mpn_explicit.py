"""TEST INFRASTRUCTURE ONLY -- numpy float64 forward + HAND-DERIVED backward of the
reference encoder (chemprop/models/mpn.py:92-171, ReLU, dropout 0, bond messages).

Purpose: pin the gradient formulae the CUDA backward implements
against torch autograd of the reference op sequence (oracle/mpn_oracle.py) -- see
tests/test_oracle_explicit.py.  CSR form, no padded [N_a,K,h] intermediate: this is the
algorithmic shape of the device path, written as slow, obviously-correct loops.
"""
import numpy as np

from .mpn_oracle import csr_from_a2b


def _aggregate_loops(M, w_bonds, in_ptr, in_idx):
    """A[a] = sum_{b in in(a)} w_b * M[b]  (mpn.py:112-118 without padded slots): the obviously-correct form."""
    A = np.zeros((len(in_ptr) - 1, M.shape[1]), dtype=M.dtype)
    for a in range(len(in_ptr) - 1):
        for j in range(in_ptr[a], in_ptr[a + 1]):
            b = in_idx[j]
            A[a] += w_bonds[b] * M[b]
    return A


def _aggregate(M, w_bonds, in_ptr, in_idx):
    """Same sums as _aggregate_loops, as one sparse [N_a, N_b] times dense product (float64; list order differs
    from the loops only in rounding at 1e-16): lets the full-size configurations (271 k bonds) run in seconds.
    tests/test_oracle_golden.py holds the two forms together."""
    if len(in_idx) < 4096:
        return _aggregate_loops(M, w_bonds, in_ptr, in_idx)
    from scipy.sparse import csr_matrix
    idx = np.asarray(in_idx, np.int64)
    w = np.ones(len(idx), M.dtype) if w_bonds is None else np.asarray(w_bonds, M.dtype)[idx]
    sp = csr_matrix((w, idx, np.asarray(in_ptr, np.int64)), shape=(len(in_ptr) - 1, M.shape[0]))
    return np.asarray(sp @ M)


def tf32_round(x):
    """Round-to-nearest (ties away) of fp32 values to the 10-bit TF32 mantissa -- what the CUDA
    path's producers apply (cvt.rna.tf32.f32) before tcgen05.mma kind::tf32."""
    u = np.asarray(x, np.float32).view(np.uint32).astype(np.uint64)
    u = ((u + 0x1000) & 0xFFFFE000).astype(np.uint32)
    return u.view(np.float32).astype(np.float64)


def mm_tf32(a, b):
    """Matrix product with both operands rounded to TF32 and exact (f64) accumulation: the
    error model of the tensor-core path, used to attribute its tolerance."""
    return tf32_round(a) @ tf32_round(b)


def bf16_round(x):
    """Round-to-nearest-even of fp32 values to bfloat16 (8 significant bits): what a kind::f16 path with bf16 operands,
    or bf16 storage of the activations, would apply."""
    u = np.asarray(x, np.float32).view(np.uint32).astype(np.uint64)
    u = ((u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000).astype(np.uint32)
    return u.view(np.float32).astype(np.float64)


def mm_bf16(a, b):
    """Matrix product with both operands rounded to bfloat16 and exact accumulation: the error model of a single-pass
    BF16 tensor-core mode (not built: tests/test_oracle_golden.py::test_bf16_error_model_misses_the_bound)."""
    return bf16_round(a) @ bf16_round(b)


def forward_backward(components, weights, d_out, depth, undirected=False,
                     aggregation="mean", aggregation_norm=100.0, mm=np.matmul):
    f_atoms, f_bonds, w_atoms, w_bonds, a2b, b2a, b2revb, a_scope, _bs, dop = components
    f64 = np.float64
    f_atoms, f_bonds = np.asarray(f_atoms, f64), np.asarray(f_bonds, f64)
    w_atoms, w_bonds = np.asarray(w_atoms, f64), np.asarray(w_bonds, f64)
    a2b, b2a, b2revb = np.asarray(a2b), np.asarray(b2a), np.asarray(b2revb)
    Wi, Wh, Wo = (np.asarray(weights[k], f64) for k in ("W_i.weight", "W_h.weight", "W_o.weight"))
    bo = np.asarray(weights["W_o.bias"], f64)
    bi = np.asarray(weights["W_i.bias"], f64) if "W_i.bias" in weights else None
    bh = np.asarray(weights["W_h.bias"], f64) if "W_h.bias" in weights else None
    d_out = np.asarray(d_out, f64)
    in_ptr, in_idx, out_ptr, out_idx, dst = csr_from_a2b(a2b.tolist(), b2a.tolist(), b2revb.tolist())
    Fa, h, G = f_atoms.shape[1], Wh.shape[0], len(a_scope)

    # ---- forward ----
    I = mm(f_bonds, Wi.T) + (bi if bi is not None else 0.0)
    Ms = [np.maximum(I, 0.0)]          # M_0 .. M_{d-1}
    Us, Ss = [], []                    # U_t = (undirected-symmetrised) input of step t, S_t
    for _ in range(depth - 1):
        U = Ms[-1]
        if undirected:
            U = (U + U[b2revb]) / 2
        A = _aggregate(U, w_bonds, in_ptr, in_idx)
        S = A[b2a] - U[b2revb]
        Us.append(U)
        Ss.append(S)
        Ms.append(np.maximum(I + mm(S, Wh.T) + (bh if bh is not None else 0.0), 0.0))
    A_d = _aggregate(Ms[-1], w_bonds, in_ptr, in_idx)
    X = np.concatenate([f_atoms, A_d], axis=1)
    H = np.maximum(mm(X, Wo.T) + bo, 0.0)
    out = np.zeros((G, h), f64)
    scale = np.zeros(len(w_atoms), f64)   # d out[g] / d (w-weighted H[a]) incl. w_atoms
    for g, (s, n) in enumerate(a_scope):
        if n == 0:
            continue
        w = w_atoms[s:s + n]
        den = w.sum() if aggregation == "mean" else (1.0 if aggregation == "sum" else aggregation_norm)
        out[g] = float(dop[g]) * (w[:, None] * H[s:s + n]).sum(0) / den
        scale[s:s + n] = w * float(dop[g]) / den

    # ---- backward ----
    mol_of = np.full(len(w_atoms), -1)
    for g, (s, n) in enumerate(a_scope):
        mol_of[s:s + n] = g
    dH = np.where(mol_of[:, None] >= 0, scale[:, None] * d_out[np.maximum(mol_of, 0)], 0.0)
    dZo = dH * (H > 0)
    grads = {"W_o.weight": mm(dZo.T, X), "W_o.bias": dZo.sum(0)}
    dA = mm(dZo, Wo[:, Fa:])
    dM = w_bonds[:, None] * dA[dst]                 # transpose of the final aggregate
    dM[0] = 0.0
    dI = np.zeros_like(I)
    dWh = np.zeros_like(Wh)
    dbh = np.zeros(h, f64)
    for t in range(depth - 1, 0, -1):
        dZ = dM * (Ms[t] > 0)
        dI += dZ
        dWh += mm(dZ.T, Ss[t - 1])
        dbh += dZ.sum(0)
        dS = mm(dZ, Wh)
        dA = _aggregate(dS, None, out_ptr, out_idx) if len(out_idx) >= 4096 else None
        if dA is None:
            dA = np.zeros((len(w_atoms), h), f64)   # reverse CSR: out-edges of each atom
            for a in range(len(w_atoms)):
                for j in range(out_ptr[a], out_ptr[a + 1]):
                    dA[a] += dS[out_idx[j]]
        # row 0 of dS scatters into A[b2a[0]] = A[0] in autograd, whose own gradient path is
        # weight-0 only; it never reaches a real bond.
        dU = w_bonds[:, None] * dA[dst] - dS[b2revb]
        dU[0] = -dS[0]                              # rev(0) = 0, w_bonds[0] = 0
        if undirected:
            dU = (dU + dU[b2revb]) / 2
        dM = dU
    dZ0 = dM * (I > 0)
    dI += dZ0
    grads["W_i.weight"] = mm(dI.T, f_bonds)
    grads["W_h.weight"] = dWh
    if bi is not None:
        grads["W_i.bias"] = dI.sum(0)
    if bh is not None:
        grads["W_h.bias"] = dbh
    return out, grads
