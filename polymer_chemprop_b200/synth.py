"""Seeded synthetic graph generators with the reference's tensor contract.

The copolymer dataset and RDKit are unavailable, so the
BASELINE configs run on graphs shaped like wD-MPNN copolymer inputs: two monomers,
four attachment atoms, stochastic inter-monomer edges including self-loop pairs
(featurization.py:575-629), monomer-fraction atom weights and
``degree_of_polym = 1 + log10(Xn)`` (featurization.py:339-364).  Bond numbering,
a2b append order, b2a and b2revb follow featurization.py:467-480 / 616-629.
"""
from typing import List

import numpy as np

from .graph import ATOM_FDIM, BOND_ONLY_FDIM, MolGraphData

_ATOM_BLOCKS = (101, 7, 6, 5, 6, 6)  # one-hot widths, featurization.py:21-43


def _atom_features(rng, n, dense):
    if dense:
        return rng.random((n, ATOM_FDIM), dtype=np.float32)
    f = np.zeros((n, ATOM_FDIM), dtype=np.float32)
    off = 0
    for w in _ATOM_BLOCKS:
        idx = rng.integers(0, min(w, 12), size=n)
        f[np.arange(n), off + idx] = 1.0
        off += w
    f[:, off] = rng.integers(0, 2, size=n)                  # aromatic
    f[:, off + 1] = rng.uniform(0.01, 0.35, size=n)         # mass * 0.01
    return f


def _bond_features(rng, n, dense):
    if dense:
        return rng.random((n, BOND_ONLY_FDIM), dtype=np.float32)
    f = np.zeros((n, BOND_ONLY_FDIM), dtype=np.float32)     # [null, 4 types, conj, ring, 7 stereo]
    f[np.arange(n), 1 + rng.integers(0, 4, size=n)] = 1.0
    f[:, 5] = rng.integers(0, 2, size=n)
    f[:, 6] = rng.integers(0, 2, size=n)
    f[np.arange(n), 7 + rng.integers(0, 7, size=n)] = 1.0
    return f


def _random_bonded_pairs(rng, n, max_deg, ring_choices, ring_p, offset=0):
    """Random tree (parent among the previous 4 atoms, degree-capped) + ring closures.
    Returns a sorted list of (a1 < a2) pairs in local numbering shifted by ``offset``."""
    deg = np.zeros(n, dtype=np.int64)
    pairs = set()
    for a in range(1, n):
        lo = max(0, a - 4)
        cand = [p for p in range(lo, a) if deg[p] < max_deg]
        if not cand:
            cand = [p for p in range(0, a) if deg[p] < max_deg] or [a - 1]
        p = int(cand[rng.integers(0, len(cand))])
        pairs.add((p, a))
        deg[p] += 1
        deg[a] += 1
    n_ring = int(rng.choice(ring_choices, p=ring_p))
    for _ in range(n_ring):
        for _try in range(8):
            if n < 4:
                break
            a = int(rng.integers(0, n))
            b = int(rng.integers(0, n))
            if a > b:
                a, b = b, a
            if b - a < 2 or (a, b) in pairs or deg[a] >= max_deg or deg[b] >= max_deg:
                continue
            pairs.add((a, b))
            deg[a] += 1
            deg[b] += 1
            break
    return sorted((a + offset, b + offset) for a, b in pairs)


def _assemble(n_atoms, pairs, pair_w, f_atoms, f_bond14, w_atoms, dop):
    """pairs: ordered list of (a1, a2) (a1 may equal a2); each yields two directed bonds
    b1 = a1->a2, b2 = a2->a1 exactly as featurization.py:467-480."""
    a2b = [[] for _ in range(n_atoms)]
    nb = 2 * len(pairs)
    b2a = np.zeros(nb, dtype=np.int64)
    b2revb = np.zeros(nb, dtype=np.int64)
    w_bonds = np.zeros(nb, dtype=np.float32)
    f_bonds = np.zeros((nb, f_atoms.shape[1] + f_bond14.shape[1]), dtype=np.float32)
    for i, (a1, a2) in enumerate(pairs):
        b1, b2 = 2 * i, 2 * i + 1
        a2b[a2].append(b1)
        b2a[b1] = a1
        a2b[a1].append(b2)
        b2a[b2] = a2
        b2revb[b1] = b2
        b2revb[b2] = b1
        w_bonds[b1], w_bonds[b2] = pair_w[i]
        f_bonds[b1, :f_atoms.shape[1]] = f_atoms[a1]
        f_bonds[b2, :f_atoms.shape[1]] = f_atoms[a2]
        f_bonds[b1, f_atoms.shape[1]:] = f_bond14[i]
        f_bonds[b2, f_atoms.shape[1]:] = f_bond14[i]
    return MolGraphData(f_atoms, f_bonds, w_atoms, w_bonds, a2b, b2a, b2revb, dop)


def copolymer_graph(rng, dense_features=False, random_edge_weights=False) -> MolGraphData:
    nA = int(rng.integers(10, 19))
    nB = int(rng.integers(6, 13))
    n = nA + nB
    pairs = _random_bonded_pairs(rng, nA, 3, [0, 1, 2], [1 / 3, 1 / 3, 1 / 3])
    pairs += _random_bonded_pairs(rng, nB, 3, [0, 1, 2], [1 / 3, 1 / 3, 1 / 3], offset=nA)
    pair_w = [(1.0, 1.0)] * len(pairs)
    # four attachment atoms: two per monomer
    att = [int(x) for x in rng.choice(nA, size=2, replace=False)] + \
          [int(nA + x) for x in rng.choice(nB, size=2, replace=False)]
    for i in range(4):
        for j in range(i, 4):                     # 10 unordered pairs incl. the 4 self pairs
            pairs.append((att[i], att[j]))
            if random_edge_weights:
                pair_w.append((float(rng.uniform(0.05, 1.0)), float(rng.uniform(0.05, 1.0))))
            else:
                pair_w.append((0.25, 0.25))
    fa = _atom_features(rng, n, dense_features)
    fb = _bond_features(rng, len(pairs), dense_features)
    fr = [(0.5, 0.5), (0.25, 0.75), (0.75, 0.25)][int(rng.integers(0, 3))]
    w_atoms = np.concatenate([np.full(nA, fr[0], np.float32), np.full(nB, fr[1], np.float32)])
    if rng.random() < 0.5:
        dop = 1
    else:
        xn = 10.0 ** rng.uniform(1.0, 3.0)
        dop = 1.0 + np.log10(xn)                  # featurization.py:364 (np.float64)
    return _assemble(n, pairs, pair_w, fa, fb, w_atoms, dop)


def qm9_graph(rng, dense_features=False) -> MolGraphData:
    n = int(rng.choice([7, 8, 9], p=[0.1, 0.2, 0.7]))
    pairs = _random_bonded_pairs(rng, n, 4, [0, 1, 2], [0.3, 0.4, 0.3])
    fa = _atom_features(rng, n, dense_features)
    fb = _bond_features(rng, len(pairs), dense_features)
    return _assemble(n, pairs, [(1.0, 1.0)] * len(pairs), fa, fb, np.ones(n, np.float32), 1)


def single_atom_graph(rng, dense_features=True) -> MolGraphData:
    """One heavy atom, no bonds (the K==1 / max(1, ...) corner, featurization.py:802-803)."""
    fa = _atom_features(rng, 1, dense_features)
    return _assemble(1, [], [], fa, np.zeros((0, BOND_ONLY_FDIM), np.float32), np.ones(1, np.float32), 1)


def empty_graph() -> MolGraphData:
    """Zero atoms: exercises the a_size == 0 readout branch (mpn.py:148-149)."""
    return MolGraphData(np.zeros((0, ATOM_FDIM), np.float32), np.zeros((0, ATOM_FDIM + BOND_ONLY_FDIM), np.float32),
                        [], [], [], [], [], 1)


def make_graphs(kind: str, n: int, seed: int = 0, dense_features: bool = False,
                random_edge_weights: bool = False) -> List[MolGraphData]:
    rng = np.random.default_rng(seed)
    if kind == "copolymer":
        return [copolymer_graph(rng, dense_features, random_edge_weights) for _ in range(n)]
    if kind == "qm9":
        return [qm9_graph(rng, dense_features) for _ in range(n)]
    raise ValueError(f"unknown synthetic graph kind {kind!r}")
