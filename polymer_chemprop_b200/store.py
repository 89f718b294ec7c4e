"""Device-resident store of featurised graphs + on-device batch assembly.

The reference keeps every featurised molecule as a host ``MolGraph`` (cache ``SMILES_TO_GRAPH``,
chemprop/data/data.py:30-40), builds each batch on the host with Python list appends
(``MoleculeDataset.batch_graph``, data.py:277-314 -> ``BatchMolGraph.__init__``, featurization.py:757-813) and
uploads it on every encoder call (chemprop/models/mpn.py:84-87).  ``DeviceGraphStore`` uploads the whole
dataset ONCE in compact form (atom rows, bond-only tails, local indices: ~6 KB per copolymer graph) and
``batch(indices)`` assembles one step's batch in HBM (``wd_store_assemble``): only the selected indices and
their start offsets cross PCIe.  The result is bit-identical to ``DeviceGraph(BatchMolGraph([graphs[i] ...]))``
(tests/test_gpu_store.py) and is consumed by ``MPN.forward`` like any other batch object.
"""
import ctypes as C
from typing import List, Sequence

import numpy as np
import torch

from . import _capi
from .graph import BatchMolGraph, MolGraphData
from .mpn import DeviceGraph


class StoreBatch:
    """One assembled batch: duck-types a BatchMolGraph for ``MPN.forward`` (its device form is already cached on it);
    ``get_components()`` materialises the host tensors from the store's host copies (tests / oracle only)."""

    def __init__(self, store: "DeviceGraphStore", indices: np.ndarray, dg: DeviceGraph):
        self._store, self.indices = store, indices
        self._wd_cache = {(dg.device.type, dg.device.index): dg}
        self.n_atoms, self.n_bonds = dg.n_atoms, dg.n_bonds
        self.atom_fdim, self.bond_fdim = dg.atom_fdim, dg.bond_fdim

    def _scope(self, sizes):
        starts = 1 + np.concatenate([[0], np.cumsum(sizes)[:-1]])
        return [(int(s), int(n)) for s, n in zip(starts, sizes)]

    @property
    def a_scope(self):          # featurization.py:795 (built on demand: a Python list of G tuples costs ~1 ms at G = 4096)
        return self._scope(self._store.n_atoms_of[self.indices])

    @property
    def b_scope(self):          # featurization.py:796
        return self._scope(self._store.n_bonds_of[self.indices])

    def to_batch_mol_graph(self) -> BatchMolGraph:
        return BatchMolGraph([self._store.graphs[int(i)] for i in self.indices])

    def get_components(self, atom_messages: bool = False):
        return self.to_batch_mol_graph().get_components(atom_messages)


class DeviceGraphStore:
    @staticmethod
    def pack(graphs: Sequence[MolGraphData]) -> dict:
        """Host-side compact form of the store (numpy, local numbering): what __init__ uploads."""
        if not graphs:
            raise ValueError("empty store")
        bad = [i for i, g in enumerate(graphs) if not g.bond_rows_start_with_atom_row]
        if bad:
            raise ValueError(f"graphs {bad[:5]}... do not build their bond rows as f_atoms[a1] + f_bond "
                             "(featurization.py:467-468): the compact store cannot hold them")
        g0 = next((g for g in graphs if g.n_bonds), graphs[0])
        atom_fdim = next((g.f_atoms.shape[1] for g in graphs if g.n_atoms), g0.f_atoms.shape[1])
        bond_fdim = g0.f_bonds.shape[1]
        tail_fdim = bond_fdim - atom_fdim
        na = np.fromiter((g.n_atoms for g in graphs), dtype=np.int64, count=len(graphs))
        nb = np.fromiter((g.n_bonds for g in graphs), dtype=np.int64, count=len(graphs))
        atom_off = np.concatenate([[0], np.cumsum(na)])
        bond_off = np.concatenate([[0], np.cumsum(nb)])
        ta, tb = int(atom_off[-1]), int(bond_off[-1])
        f_atoms = np.zeros((max(ta, 1), atom_fdim), np.float32)
        tail = np.zeros((max(tb, 1), max(tail_fdim, 1)), np.float32)
        w_atoms, w_bonds = np.zeros(max(ta, 1), np.float32), np.zeros(max(tb, 1), np.float32)
        b2a, b2revb = np.zeros(max(tb, 1), np.int32), np.zeros(max(tb, 1), np.int32)
        in_idx, out_idx = np.zeros(max(tb, 1), np.int32), np.zeros(max(tb, 1), np.int32)
        in_ptr_local = np.zeros(max(ta, 1), np.int32)
        max_in_degree, adjacent = 0, True
        for g, a0, b0 in zip(graphs, atom_off[:-1], bond_off[:-1]):
            if g.n_atoms:
                f_atoms[a0:a0 + g.n_atoms] = g.f_atoms
                w_atoms[a0:a0 + g.n_atoms] = g.w_atoms
                deg = np.fromiter((len(r) for r in g.a2b), dtype=np.int64, count=g.n_atoms)
                in_ptr_local[a0:a0 + g.n_atoms] = np.concatenate([[0], np.cumsum(deg)[:-1]])
                max_in_degree = max(max_in_degree, int(deg.max()))
            if g.n_bonds:
                tail[b0:b0 + g.n_bonds, :tail_fdim] = g.f_bonds[:, atom_fdim:]
                w_bonds[b0:b0 + g.n_bonds] = g.w_bonds
                b2a[b0:b0 + g.n_bonds] = g.b2a
                b2revb[b0:b0 + g.n_bonds] = g.b2revb
                ins = np.fromiter((b for r in g.a2b for b in r), dtype=np.int64, count=-1)
                if ins.size != g.n_bonds or np.unique(ins).size != g.n_bonds:
                    raise ValueError("every bond must be listed in exactly one a2b row (featurization.py:471-476)")
                in_idx[b0:b0 + g.n_bonds] = ins
                # out-edges of atom a = bonds with b2a == a, ascending bond id (what wd_csr_build produces)
                out_idx[b0:b0 + g.n_bonds] = np.argsort(g.b2a, kind="stable")
                # the gathered depth step's TMA path needs adjacent reverse pairs and an even bond count per graph
                adjacent &= g.n_bonds % 2 == 0 and bool(np.array_equal(g.b2revb, np.arange(g.n_bonds) ^ 1))
                # out-degree == in-degree per atom is what lets the reverse CSR share in_ptr (wd_csr_build checks dst)
                if not np.array_equal(np.bincount(g.b2a, minlength=g.n_atoms), deg):
                    raise ValueError("out-degree != in-degree for some atom: not a graph of reverse bond pairs")
        dop = np.fromiter((float(g.degree_of_polym) for g in graphs), dtype=np.float32, count=len(graphs))

        return dict(atom_fdim=atom_fdim, bond_fdim=bond_fdim, tail_fdim=tail_fdim, n_atoms_of=na, n_bonds_of=nb,
                    max_in_degree=max_in_degree, rev_adjacent=bool(adjacent), f_atoms=f_atoms, bond_tail=tail, w_atoms=w_atoms,
                    w_bonds=w_bonds, b2a=b2a, b2revb=b2revb, in_idx=in_idx, out_idx=out_idx, in_ptr_local=in_ptr_local,
                    atom_off=atom_off.astype(np.int64), bond_off=bond_off.astype(np.int64), dop=dop)

    @staticmethod
    def _tile(pk: dict, k: int) -> dict:
        """The packed store repeated k times (graphs are held in LOCAL numbering, so their arrays tile as they are;
        only the per-graph offsets continue).  Used to stage datasets of benchmark size -- e.g. BASELINE.json's 10^6
        graphs -- from a few thousand distinct synthetic graphs without packing each copy in Python."""
        out = dict(pk)
        ta, tb = int(pk["atom_off"][-1]), int(pk["bond_off"][-1])
        for key, rows in (("f_atoms", ta), ("w_atoms", ta), ("in_ptr_local", ta), ("bond_tail", tb), ("w_bonds", tb),
                          ("b2a", tb), ("b2revb", tb), ("in_idx", tb), ("out_idx", tb)):
            x = pk[key][:rows]
            out[key] = np.concatenate([x] * k, axis=0) if rows else pk[key]
        for key in ("n_atoms_of", "n_bonds_of", "dop"):
            out[key] = np.tile(pk[key], k)
        out["atom_off"] = np.concatenate([[0], np.cumsum(out["n_atoms_of"])]).astype(np.int64)
        out["bond_off"] = np.concatenate([[0], np.cumsum(out["n_bonds_of"])]).astype(np.int64)
        return out

    def __init__(self, graphs: Sequence[MolGraphData], device: torch.device, repeat: int = 1):
        """repeat > 1: the store holds ``repeat`` copies of ``graphs`` back to back (index i is graphs[i % len(graphs)])."""
        if device.type != "cuda":
            raise RuntimeError("DeviceGraphStore lives in GPU memory (no CPU fallback)")
        base = list(graphs)
        self.device = device
        pk = self.pack(base)
        if repeat > 1:
            pk = self._tile(pk, repeat)
        self.graphs: List[MolGraphData] = base * max(1, repeat)
        self.atom_fdim, self.bond_fdim, self.tail_fdim = pk["atom_fdim"], pk["bond_fdim"], pk["tail_fdim"]
        self.n_atoms_of, self.n_bonds_of = pk["n_atoms_of"], pk["n_bonds_of"]
        self.max_in_degree, self.rev_adjacent = pk["max_in_degree"], pk["rev_adjacent"]

        def up(x):
            return torch.from_numpy(np.ascontiguousarray(x)).to(device)

        self._t = {k: up(pk[k]) for k in ("f_atoms", "bond_tail", "w_atoms", "w_bonds", "b2a", "b2revb", "in_idx", "out_idx",
                                          "in_ptr_local", "atom_off", "bond_off", "dop")}
        p = _capi.ptr
        self.struct = _capi.WdStore(len(self.graphs), self.atom_fdim, self.tail_fdim,
                                    *[p(self._t[k]) for k in ("f_atoms", "bond_tail", "w_atoms", "w_bonds", "b2a", "b2revb",
                                                              "in_idx", "out_idx", "in_ptr_local", "atom_off", "bond_off",
                                                              "dop")])
        self.nbytes = sum(t.numel() * t.element_size() for t in self._t.values())
        self._staging = []          # ring of pinned staging buffers [tensor, event of its last upload]
        self._staging_next = 0

    def _stage(self, nbytes: int):
        """A pinned host buffer of >= nbytes from a small ring (a fresh cudaHostAlloc per batch costs ~1 ms)."""
        if len(self._staging) < 4:
            self._staging.append([torch.empty(max(nbytes, 1 << 16), dtype=torch.uint8).pin_memory(), None])
            slot = self._staging[-1]
        else:
            slot = self._staging[self._staging_next % 4]
            self._staging_next += 1
            if slot[1] is not None:
                slot[1].synchronize()                       # its previous upload finished long ago
            if slot[0].numel() < nbytes:
                slot[0] = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
        return slot

    def __len__(self):
        return len(self.graphs)

    def batch(self, indices, stream=None) -> StoreBatch:
        """Assembles the batch of graphs ``indices`` (store order positions, repeats allowed) on the device.
        ``stream``: assemble on that CUDA stream (e.g. a prefetch stream); the consumer waits on the batch's event."""
        idx = np.ascontiguousarray(np.asarray(indices, dtype=np.int64).reshape(-1))
        if idx.size == 0 or idx.min() < 0 or idx.max() >= len(self.graphs):
            raise ValueError("indices out of range or empty")
        with _capi.device_guard(self.device):    # the library launches on the current device
            if stream is not None:
                with torch.cuda.stream(stream):
                    dg = self._assemble(idx)
            else:
                dg = self._assemble(idx)
        return StoreBatch(self, idx, dg)

    def _assemble(self, idx: np.ndarray) -> DeviceGraph:
        dev, G = self.device, idx.size
        na, nb = self.n_atoms_of[idx], self.n_bonds_of[idx]
        a_start = 1 + np.concatenate([[0], np.cumsum(na)[:-1]])        # featurization.py:795-800
        b_start = 1 + np.concatenate([[0], np.cumsum(nb)[:-1]])
        n_atoms, n_bonds = int(1 + na.sum()), int(1 + nb.sum())
        if n_bonds >= 2 ** 31 or n_atoms >= 2 ** 31:
            raise NotImplementedError("batch too large for int32 indices")
        # one pinned staging buffer, one H2D copy: [sel int64 | a_start, a_size, b_start, b_size int32]
        slot = self._stage(24 * G)
        host = slot[0][:24 * G]
        hv = host.numpy()
        hv[:8 * G].view(np.int64)[:] = idx
        ints = hv[8 * G:].view(np.int32).reshape(4, G)
        ints[0], ints[1], ints[2], ints[3] = a_start, na, b_start, nb
        staged = host.to(dev, non_blocking=True)
        slot[1] = torch.cuda.Event()
        slot[1].record(torch.cuda.current_stream(dev))
        sel = staged[:8 * G].view(torch.int64)
        scopes = staged[8 * G:].view(torch.int32).view(4, G)
        dg = DeviceGraph.__new__(DeviceGraph)
        dg.device, dg.compact_upload = dev, True
        dg.atom_fdim, dg.bond_fdim = self.atom_fdim, self.bond_fdim
        dg.n_atoms, dg.n_bonds, dg.n_mols = n_atoms, n_bonds, G
        ld_a, ld_b = (self.atom_fdim + 3) // 4 * 4, (self.bond_fdim + 3) // 4 * 4
        f32, i32 = dict(device=dev, dtype=torch.float32), dict(device=dev, dtype=torch.int32)
        H, sid = _capi.empty_high_water, id(self)        # capacities: largest bucket seen so far (see _capi)
        dg.f_atoms = H((sid, "fa"), (n_atoms, ld_a), torch.float32, dev)
        dg.f_bonds = H((sid, "fb"), (n_bonds, ld_b), torch.float32, dev)
        dg.w_atoms, dg.w_bonds = H((sid, "wa"), (n_atoms,), torch.float32, dev), H((sid, "wb"), (n_bonds,), torch.float32, dev)
        dg.in_ptr = H((sid, "ip"), (n_atoms + 1,), torch.int32, dev)
        dg.in_idx = H((sid, "ii"), (max(n_bonds - 1, 1),), torch.int32, dev)
        dg.out_idx = H((sid, "oi"), (max(n_bonds - 1, 1),), torch.int32, dev)
        dg.src, dg.rev, dg.dst = (H((sid, k), (n_bonds,), torch.int32, dev) for k in ("src", "rev", "dst"))
        dg.mol_of = H((sid, "mo"), (n_atoms,), torch.int32, dev)
        dg.dop = torch.empty(G, **f32)
        dg.a_start, dg.a_size = scopes[0], scopes[1]
        dg.flags = torch.zeros(4, **i32)
        bond_mol = H((sid, "bm"), (n_bonds,), torch.int32, dev)
        p = _capi.ptr
        dg.struct = _capi.WdGraph(n_atoms, n_bonds, G, p(dg.f_atoms), p(dg.f_bonds), p(dg.w_atoms), p(dg.w_bonds),
                                  p(dg.in_ptr), p(dg.in_idx), p(dg.out_idx), p(dg.src), p(dg.rev), p(dg.dst), p(dg.mol_of),
                                  p(dg.a_start), p(dg.a_size), p(dg.dop), ld_a, ld_b, int(self.rev_adjacent))
        _capi.check(_capi.lib().wd_store_assemble(C.byref(self.struct), p(sel), p(scopes[0]), p(scopes[1]), p(scopes[2]),
                                                  p(scopes[3]), G, C.byref(dg.struct), p(bond_mol),
                                                  _capi.current_stream(dev)))
        dg._store_keepalive = (staged, bond_mol, self)           # read by the kernels still in flight
        dg.build_stream = torch.cuda.current_stream(dev)
        dg.ready = torch.cuda.Event()
        dg.ready.record(dg.build_stream)
        dg.rev_adjacent, dg.violations, dg.max_in_degree = self.rev_adjacent, 0, self.max_in_degree
        dg.flags_host, dg._validate, dg._checked = None, False, True       # the store validated its graphs once
        dg._int64_keepalive = dg._host_keepalive = dg._dense_keepalive = None
        return dg
