"""B200-native wD-MPNN encoder hot path (drop-in for chemprop.models.mpn.MPN)."""
from .graph import BatchMolGraph, MolGraphData, ATOM_FDIM, BOND_FDIM  # noqa: F401
from .mpn import MPN, MPNEncoder, DeviceGraph, install  # noqa: F401,E402
from .store import DeviceGraphStore, StoreBatch  # noqa: F401,E402
from .optim import FlatAdam, NoamLR  # noqa: F401,E402
from .head import ffn_mse_loss, ffn_predict  # noqa: F401,E402
