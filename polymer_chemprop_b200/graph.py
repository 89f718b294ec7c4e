"""Host-side mirror of the reference's graph containers for the encoder hot path.

``MolGraphData`` holds one featurised molecule / polymer (the fields
``BatchMolGraph.__init__`` reads from a ``MolGraph``:
chemprop/features/featurization.py:757-813).  ``BatchMolGraph`` is the
drop-in batch container: same attribute names, same row-0 padding, same
``get_components()`` 10-tuple (featurization.py:815-846), but assembled with
vectorised numpy instead of Python list appends.

RDKit featurisation (SMILES -> MolGraph) is out of scope;
graphs reach this module already featurised.
"""
from typing import List, Sequence, Tuple

import numpy as np
import torch

ATOM_FDIM = 133   # featurization.py:21-43 (101+7+6+5+6+6+1+1)
BOND_ONLY_FDIM = 14  # featurization.py:229-250
BOND_FDIM = ATOM_FDIM + BOND_ONLY_FDIM


class MolGraphData:
    """One molecule in the reference's local numbering (atoms/bonds start at 0).

    a2b[a] lists incoming directed bonds of atom ``a`` in append order
    (featurization.py:467-480); ``b2a[b]`` is the source atom; ``b2revb[b]`` the
    reverse bond.
    """

    __slots__ = ("n_atoms", "n_bonds", "f_atoms", "f_bonds", "w_atoms", "w_bonds", "a2b", "b2a",
                 "b2revb", "degree_of_polym", "overwrite_default_atom_features",
                 "overwrite_default_bond_features", "_prefix_ok")

    def __init__(self, f_atoms, f_bonds, w_atoms, w_bonds, a2b, b2a, b2revb, degree_of_polym=1):
        self.f_atoms = np.asarray(f_atoms, dtype=np.float32)
        self.f_bonds = np.asarray(f_bonds, dtype=np.float32)
        if self.f_atoms.ndim != 2:
            self.f_atoms = self.f_atoms.reshape(len(w_atoms), ATOM_FDIM if len(w_atoms) == 0 else -1)
        if self.f_bonds.ndim != 2:
            self.f_bonds = self.f_bonds.reshape(len(w_bonds), BOND_FDIM if len(w_bonds) == 0 else -1)
        self.w_atoms = np.asarray(w_atoms, dtype=np.float32)
        self.w_bonds = np.asarray(w_bonds, dtype=np.float32)
        self.a2b = [list(map(int, r)) for r in a2b]
        self.b2a = np.asarray(b2a, dtype=np.int64)
        self.b2revb = np.asarray(b2revb, dtype=np.int64)
        self.n_atoms = len(self.w_atoms)
        self.n_bonds = len(self.w_bonds)
        self.degree_of_polym = degree_of_polym
        self.overwrite_default_atom_features = False
        self.overwrite_default_bond_features = False
        self._prefix_ok = None

    @property
    def bond_rows_start_with_atom_row(self) -> bool:
        """True iff f_bonds[b, :F_a] == f_atoms[b2a[b]] bit for bit for every bond -- how MolGraph builds
        its bond rows (featurization.py:467-468, :616-629).  Checked exactly, once per molecule."""
        if self._prefix_ok is None:
            fa = self.f_atoms.shape[1]
            if self.n_bonds == 0:
                self._prefix_ok = True
            elif self.n_atoms == 0 or self.f_bonds.shape[1] < fa:
                self._prefix_ok = False
            else:
                a = self.f_bonds[:, :fa].view(np.uint32)
                b = self.f_atoms[self.b2a].view(np.uint32)
                self._prefix_ok = bool(np.array_equal(a, b))
        return self._prefix_ok

    @staticmethod
    def of(g) -> "MolGraphData":
        """``g`` itself, or the array form of a REFERENCE ``MolGraph`` (chemprop/features/featurization.py:377-640:
        Python lists ``f_atoms, f_bonds, w_atoms, w_bonds, a2b, b2a, b2revb`` + ``degree_of_polym``).  The conversion is
        cached on the object, as the reference caches its MolGraphs in SMILES_TO_GRAPH (data/data.py:30-40), so it is
        paid once per molecule, not once per batch."""
        if isinstance(g, MolGraphData):
            return g
        cached = getattr(g, "_wd_data", None)
        if cached is None:
            cached = MolGraphData(g.f_atoms, g.f_bonds, g.w_atoms, g.w_bonds, g.a2b, g.b2a, g.b2revb,
                                  getattr(g, "degree_of_polym", 1))
            cached.overwrite_default_atom_features = getattr(g, "overwrite_default_atom_features", False)
            cached.overwrite_default_bond_features = getattr(g, "overwrite_default_bond_features", False)
            try:
                g._wd_data = cached
            except AttributeError:
                pass
        return cached

    def as_reference_lists(self):
        """Duck-typed object with plain Python lists, consumable by the *reference's*
        BatchMolGraph.__init__ (used only by oracle/ scripts and tests)."""
        import types
        return types.SimpleNamespace(
            n_atoms=self.n_atoms, n_bonds=self.n_bonds,
            f_atoms=self.f_atoms.tolist(), f_bonds=self.f_bonds.tolist(),
            w_atoms=self.w_atoms.tolist(), w_bonds=self.w_bonds.tolist(),
            a2b=[list(r) for r in self.a2b], b2a=self.b2a.tolist(), b2revb=self.b2revb.tolist(),
            degree_of_polym=self.degree_of_polym,
            overwrite_default_atom_features=False, overwrite_default_bond_features=False)


class BatchMolGraph:
    """Drop-in for chemprop.features.BatchMolGraph (featurization.py:743-875).

    Row 0 of every array is padding; scopes start at 1.  Tensors are created on
    the CPU exactly like the reference; ``to(device)`` moves them once so the
    encoder's per-call ``.to(device)`` (mpn.py:84-87) is a no-op afterwards.
    """

    def __init__(self, mol_graphs: Sequence[MolGraphData]):
        """``mol_graphs``: MolGraphData objects, or the reference's own MolGraph objects (what
        MoleculeDataset.batch_graph hands over, data/data.py:277-314, once install() has put this class in place)."""
        mol_graphs = [MolGraphData.of(g) for g in mol_graphs]
        g0 = mol_graphs[0]
        self.overwrite_default_atom_features = g0.overwrite_default_atom_features
        self.overwrite_default_bond_features = g0.overwrite_default_bond_features
        self.atom_fdim = g0.f_atoms.shape[1] if g0.n_atoms else ATOM_FDIM
        self.bond_fdim = g0.f_bonds.shape[1] if g0.n_bonds else self.atom_fdim + BOND_ONLY_FDIM

        na = np.fromiter((g.n_atoms for g in mol_graphs), dtype=np.int64, count=len(mol_graphs))
        nb = np.fromiter((g.n_bonds for g in mol_graphs), dtype=np.int64, count=len(mol_graphs))
        a_start = 1 + np.concatenate([[0], np.cumsum(na)[:-1]])
        b_start = 1 + np.concatenate([[0], np.cumsum(nb)[:-1]])
        self.n_atoms = int(1 + na.sum())
        self.n_bonds = int(1 + nb.sum())
        self.a_scope: List[Tuple[int, int]] = [(int(s), int(n)) for s, n in zip(a_start, na)]
        self.b_scope: List[Tuple[int, int]] = [(int(s), int(n)) for s, n in zip(b_start, nb)]
        self.degree_of_polym = [g.degree_of_polym for g in mol_graphs]
        # array forms of a_scope / degree_of_polym for the upload (the 10-tuple keeps the reference's Python lists;
        # converting 4096 tuples per batch costs milliseconds of host time, more than the encoder step itself)
        self.a_scope_array = torch.from_numpy(np.concatenate([a_start, na]).astype(np.int32))       # starts | sizes
        self.dop_array = torch.from_numpy(np.fromiter((float(x) for x in self.degree_of_polym), dtype=np.float32,
                                                      count=len(mol_graphs)))

        f_atoms = np.zeros((self.n_atoms, self.atom_fdim), dtype=np.float32)
        f_bonds = np.zeros((self.n_bonds, self.bond_fdim), dtype=np.float32)
        w_atoms = np.zeros(self.n_atoms, dtype=np.float32)
        w_bonds = np.zeros(self.n_bonds, dtype=np.float32)
        b2a = np.zeros(self.n_bonds, dtype=np.int64)
        b2revb = np.zeros(self.n_bonds, dtype=np.int64)
        max_nb = 1  # max(1, ...) as featurization.py:802-803
        for g in mol_graphs:
            for r in g.a2b:
                if len(r) > max_nb:
                    max_nb = len(r)
        self.max_num_bonds = max_nb
        a2b = np.zeros((self.n_atoms, max_nb), dtype=np.int64)
        for g, a0, b0 in zip(mol_graphs, a_start, b_start):
            if g.n_atoms:
                f_atoms[a0:a0 + g.n_atoms] = g.f_atoms
                w_atoms[a0:a0 + g.n_atoms] = g.w_atoms
                for i, r in enumerate(g.a2b):
                    if r:
                        a2b[a0 + i, :len(r)] = np.asarray(r, dtype=np.int64) + b0
            if g.n_bonds:
                f_bonds[b0:b0 + g.n_bonds] = g.f_bonds
                w_bonds[b0:b0 + g.n_bonds] = g.w_bonds
                b2a[b0:b0 + g.n_bonds] = g.b2a + a0
                b2revb[b0:b0 + g.n_bonds] = g.b2revb + b0

        # Compact form of f_bonds for the upload (csrc/features.cu): when every molecule's bond rows start with
        # the source atom's row -- verified above, molecule by molecule -- only the bond-only tail is new data.
        self.f_bonds_tail = None
        if self.bond_fdim > self.atom_fdim and all(g.bond_rows_start_with_atom_row for g in mol_graphs):
            self.f_bonds_tail = torch.from_numpy(np.ascontiguousarray(f_bonds[:, self.atom_fdim:]))
        self.f_atoms = torch.from_numpy(f_atoms)
        self.f_bonds = torch.from_numpy(f_bonds)
        self.w_atoms = torch.from_numpy(w_atoms)
        self.w_bonds = torch.from_numpy(w_bonds)
        self.a2b = torch.from_numpy(a2b)
        self.b2a = torch.from_numpy(b2a)
        self.b2revb = torch.from_numpy(b2revb)
        self.b2b = None
        self.a2a = None

    def get_components(self, atom_messages: bool = False):
        """Same 10-tuple, same order as featurization.py:845-846."""
        if atom_messages:
            raise NotImplementedError(
                "atom_messages=True is not accelerated by the B200 encoder (SURVEY.md section 8a); "
                "use the reference module for that mode")
        return (self.f_atoms, self.f_bonds, self.w_atoms, self.w_bonds, self.a2b, self.b2a,
                self.b2revb, self.a_scope, self.b_scope, self.degree_of_polym)

    def get_a2a(self) -> torch.Tensor:
        """featurization.py:862-875."""
        if self.a2a is None:
            self.a2a = self.b2a[self.a2b]
        return self.a2a

    def get_b2b(self) -> torch.Tensor:
        """featurization.py:848-860."""
        if self.b2b is None:
            b2b = self.a2b[self.b2a]
            revmask = (b2b != self.b2revb.unsqueeze(1)).long()
            self.b2b = b2b * revmask
        return self.b2b

    _TENSORS = ("f_atoms", "f_bonds", "w_atoms", "w_bonds", "a2b", "b2a", "b2revb", "f_bonds_tail", "a_scope_array",
                "dop_array")

    # what DeviceGraph uploads for a batch that carries the bond-only tail (compact upload), in arena order
    _UPLOAD = ("f_atoms", "f_bonds_tail", "w_atoms", "w_bonds", "a2b", "b2a", "b2revb", "a_scope_array", "dop_array")

    def pack_arena(self, pin: bool = True) -> "BatchMolGraph":
        """Lays everything the compact upload needs out in ONE host buffer (``_arena``; every array at a 256-byte
        aligned offset, recorded in ``_arena_layout``) and turns the attributes into views of it, so that DeviceGraph
        uploads the batch with a single copy.  Values and dtypes are unchanged."""
        if self.f_bonds_tail is None or getattr(self, "_arena", None) is not None or self.f_atoms.device.type != "cpu":
            return self
        parts, off = {}, 0
        for name in self._UPLOAD:
            t = getattr(self, name).contiguous()
            parts[name] = (off, t)
            off += (t.numel() * t.element_size() + 255) // 256 * 256
        arena = torch.empty(max(off, 256), dtype=torch.uint8)
        if pin:
            arena = arena.pin_memory()
        layout = {}
        for name, (o, t) in parts.items():
            nb = t.numel() * t.element_size()
            view = arena[o:o + nb].view(t.dtype).view(t.shape)
            view.copy_(t)
            setattr(self, name, view)
            layout[name] = (o, nb)
        self._arena, self._arena_layout = arena, layout
        return self

    def pin_memory(self) -> "BatchMolGraph":
        """Page-locks the host tensors in place (what a DataLoader with pin_memory=True does to its batches), so
        that DeviceGraph.prefetch can upload them asynchronously on a copy stream; a batch that carries the
        bond-only tail is packed into one pinned arena first (pack_arena)."""
        self.pack_arena(pin=True)
        for name in self._TENSORS:
            t = getattr(self, name)
            if t is not None and t.device.type == "cpu" and not t.is_pinned():
                setattr(self, name, t.pin_memory())
        return self

    def fresh_view(self) -> "BatchMolGraph":
        """A new batch object over the same host tensors with nothing cached on it (no device copy)."""
        import copy
        other = copy.copy(self)
        other.__dict__.pop("_wd_cache", None)
        return other

    def to(self, device, pin: bool = False) -> "BatchMolGraph":
        """Moves the seven tensors to ``device`` in place (extension over the reference:
        keeps batches resident so mpn.py:84-87's per-call upload disappears)."""
        for name in ("f_atoms", "f_bonds", "w_atoms", "w_bonds", "a2b", "b2a", "b2revb", "f_bonds_tail"):
            t = getattr(self, name)
            if t is None:
                continue
            if pin and t.device.type == "cpu":
                t = t.pin_memory()
            setattr(self, name, t.to(device, non_blocking=True))
        if hasattr(self, "_wd_cache"):
            del self._wd_cache
        self._arena = None
        return self
