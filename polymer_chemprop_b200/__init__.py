"""B200-native wD-MPNN encoder hot path (drop-in for chemprop.models.mpn.MPN)."""
from .graph import BatchMolGraph, MolGraphData, ATOM_FDIM, BOND_FDIM  # noqa: F401
from .mpn import MPN, MPNEncoder, DeviceGraph  # noqa: F401,E402
