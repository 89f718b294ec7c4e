"""TEST INFRASTRUCTURE ONLY -- import shim for the *unmodified* reference encoder.

Only usable where /root/reference exists (the build container, never the GPU
box).  It is used by ``oracle/make_golden.py`` to generate the committed
fixtures under ``tests/golden/`` and by ``tests/test_oracle_vs_reference.py``
(skipped when the tree is absent).  Nothing under the product package imports
this file.

Recipe:
  * ``rdkit*``, ``tensorboardX`` are replaced by MagicMock modules, ``tap`` by a
    tiny real module, because they are not installed here and the encoder never
    calls into them;
  * a synthetic parent package ``chemprop`` with ``__path__`` pointing into the
    reference tree is registered so chemprop/__init__.py (which imports the CLI
    stack) is not executed;
  * ``chemprop.data`` is imported before ``chemprop.models.mpn`` to resolve the
    args.py <-> data/utils.py import cycle.
No reference source is copied or modified.
"""
import importlib
import os
import sys
import types
from unittest import mock

REFERENCE_ROOT = os.environ.get("WDMPNN_REFERENCE_ROOT", "/root/reference")


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "chemprop", "models"))


_loaded = None


def load_reference():
    """Returns a namespace with the reference's MPN, MPNEncoder, BatchMolGraph."""
    global _loaded
    if _loaded is not None:
        return _loaded
    if not reference_available():
        raise RuntimeError(f"reference tree not found at {REFERENCE_ROOT}")
    for name in ["rdkit", "rdkit.Chem", "rdkit.Chem.AllChem", "rdkit.Chem.PandasTools",
                 "rdkit.DataStructs", "rdkit.Chem.Scaffolds", "rdkit.Chem.Scaffolds.MurckoScaffold",
                 "rdkit.Chem.rdchem", "tensorboardX"]:
        if name not in sys.modules:
            m = mock.MagicMock(name=name)
            m.__path__ = []
            m.__spec__ = None
            sys.modules[name] = m
    if "tap" not in sys.modules:
        tap = types.ModuleType("tap")

        class Tap:  # minimal stand-in for typed-argument-parser
            def __init__(self, *a, **k):
                pass

        tap.Tap = Tap
        sys.modules["tap"] = tap
    if "chemprop" not in sys.modules:
        pkg = types.ModuleType("chemprop")
        pkg.__path__ = [os.path.join(REFERENCE_ROOT, "chemprop")]
        sys.modules["chemprop"] = pkg
    importlib.import_module("chemprop.data")
    mpn = importlib.import_module("chemprop.models.mpn")
    feat = importlib.import_module("chemprop.features.featurization")
    _loaded = types.SimpleNamespace(MPN=mpn.MPN, MPNEncoder=mpn.MPNEncoder,
                                    BatchMolGraph=feat.BatchMolGraph, featurization=feat,
                                    mpn_module=mpn)
    return _loaded


def make_args(hidden_size=300, depth=3, bias=False, dropout=0.0, activation="ReLU",
              undirected=False, aggregation="mean", aggregation_norm=100,
              device="cpu", number_of_molecules=1, mpn_shared=False):
    """The 17 TrainArgs fields the encoder reads."""
    import torch
    return types.SimpleNamespace(
        hidden_size=hidden_size, depth=depth, bias=bias, dropout=dropout, activation=activation,
        atom_messages=False, undirected=undirected, aggregation=aggregation,
        aggregation_norm=aggregation_norm, atom_descriptors=None, atom_descriptors_size=0,
        device=torch.device(device), mpn_shared=mpn_shared, number_of_molecules=number_of_molecules,
        features_only=False, use_input_features=False,
        overwrite_default_atom_features=False, overwrite_default_bond_features=False)
