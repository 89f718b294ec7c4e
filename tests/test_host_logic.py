"""CPU: host-side mirror of the reference interface and the C-ABI library surface (no GPU compute)."""
import ctypes
import os
import re

import pytest
import torch

import _util
from polymer_chemprop_b200 import MPN, MPNEncoder, BatchMolGraph, _capi, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "wdmpnn.h")).read()
    declared = set(re.findall(r"\b(wd_[a-z_]+)\s*\(", header))
    assert declared == set(_capi.EXPORTS)
    lib = _capi.lib()
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.wd_abi_version() == _capi.ABI_VERSION


def test_workspace_bytes_and_error_convention():
    lib = _capi.lib()
    cfg = _capi.WdConfig(300, 133, 147, 3, 0, 0, 0, 0, 0, 100.0)
    small = lib.wd_workspace_bytes(ctypes.byref(cfg), 1152, 3314, 50)
    big = lib.wd_workspace_bytes(ctypes.byref(cfg), 94223, 270640, 4096)
    assert 0 < small < big < 8 << 30
    bad = _capi.WdConfig(300, 133, 147, 0, 0, 0, 0, 0, 0, 100.0)
    assert lib.wd_workspace_bytes(ctypes.byref(bad), 10, 10, 1) == 0
    rc = lib.wd_encoder_forward(None, None, ctypes.byref(bad), None, 0, None, None)
    assert rc == _capi.WD_ERR_INVALID_ARG
    with pytest.raises(ValueError):
        _capi.check(rc)


def test_state_dict_contract_matches_reference_names():
    """chemprop/utils.py:114-115, :194 rely on these exact keys / shapes."""
    m = MPN(_util.make_args(device="cpu"))
    sd = m.state_dict()
    assert list(sd) == ["encoder.0.cached_zero_vector", "encoder.0.W_i.weight", "encoder.0.W_h.weight",
                        "encoder.0.W_o.weight", "encoder.0.W_o.bias"]
    assert sd["encoder.0.W_i.weight"].shape == (300, 147) and sd["encoder.0.W_o.weight"].shape == (300, 433)
    assert not m.encoder[0].cached_zero_vector.requires_grad
    mb = MPN(_util.make_args(device="cpu", bias=True))
    assert "encoder.0.W_i.bias" in mb.state_dict() and "encoder.0.W_h.bias" in mb.state_dict()
    shared = MPN(_util.make_args(device="cpu", number_of_molecules=2, mpn_shared=True))
    assert shared.encoder[0] is shared.encoder[1]


def test_unsupported_modes_raise_before_any_launch():
    args = _util.make_args(device="cpu")
    args.atom_messages = True
    with pytest.raises(NotImplementedError):
        MPN(args)
    args = _util.make_args(device="cpu", activation="PReLU")
    with pytest.raises(NotImplementedError):
        MPN(args)
    args = _util.make_args(device="cpu")
    args.atom_descriptors = "descriptor"
    with pytest.raises(NotImplementedError):
        MPN(args)


def test_no_cpu_fallback():
    m = MPN(_util.make_args(device="cpu"))
    bmg = BatchMolGraph(synth.make_graphs("qm9", 2, 0))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        m([bmg])
    import sys
    if "chemprop" not in sys.modules:          # SMILES input needs the reference's RDKit featuriser
        with pytest.raises(NotImplementedError):
            m([["CCO"]])
