"""TEST INFRASTRUCTURE ONLY -- golden vectors of the reference's FULL model around the encoder.

Run in the build container (needs /root/reference):  python -m oracle.make_golden_model
Builds the UNMODIFIED chemprop.models.model.MoleculeModel (model.py:14-194: MPN encoder + FFN head,
initialize_weights) through oracle/ref_shim.py, runs one regression training step's forward and loss
backward exactly as chemprop/train/train.py:46-74 does (MSELoss(reduction='none') * mask, sum / mask.sum())
and stores the state_dict, predictions, loss and all parameter gradients in tests/golden/model/*.npz.
BASELINE.json's north star bounds tensor-core paths by 1e-4 relative on these PREDICTIONS.
"""
import importlib
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle import ref_shim  # noqa: E402
from polymer_chemprop_b200 import synth  # noqa: E402

OUT_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden", "model")

CASES = {
    # name: (kind, graphs, seed, model config)
    "model_copoly_b50": ("copolymer", 50, 0, dict(hidden_size=300, depth=3, ffn_num_layers=2, ffn_hidden_size=300, num_tasks=1)),
    "model_copoly_b12_tasks3": ("copolymer", 12, 5, dict(hidden_size=64, depth=4, ffn_num_layers=3, ffn_hidden_size=48, num_tasks=3,
                                                         bias=True)),
}


def model_args(hidden_size, depth, ffn_num_layers, ffn_hidden_size, num_tasks, **enc):
    a = ref_shim.make_args(hidden_size=hidden_size, depth=depth, **enc)
    a.dataset_type, a.num_tasks, a.features_size = "regression", num_tasks, 0
    a.ffn_num_layers, a.ffn_hidden_size = ffn_num_layers, ffn_hidden_size
    a.checkpoint_frzn, a.freeze_first_only, a.frzn_ffn_layers = None, False, 0
    a.multiclass_num_classes, a.spectra_activation = 3, None
    return a


def run_reference(kind, n, seed, cfg):
    ref = ref_shim.load_reference()
    MoleculeModel = importlib.import_module("chemprop.models.model").MoleculeModel
    graphs = synth.make_graphs(kind, n, seed, random_edge_weights=(seed % 2 == 1))
    bmg = ref.BatchMolGraph([g.as_reference_lists() for g in graphs])
    torch.manual_seed(seed)
    model = MoleculeModel(model_args(**cfg))                    # initialize_weights inside (nn_utils.py:102-112)
    with torch.no_grad():
        model.encoder.encoder[0].W_o.bias.normal_(0.0, 0.1)     # exercise the biases too
    model.train()
    rng = np.random.default_rng(seed + 100)
    targets = torch.from_numpy(rng.normal(size=(n, cfg["num_tasks"])).astype(np.float32))
    mask = torch.from_numpy(rng.random((n, cfg["num_tasks"])) > 0.1)    # ~10 % missing targets (train.py:46-47)
    model.zero_grad()
    preds = model([bmg])
    loss = torch.nn.MSELoss(reduction="none")(preds, targets) * mask     # train.py:70 (weights of 1)
    loss = loss.sum() / mask.sum()                                       # :71
    loss.backward()
    sd = {k: v.detach().numpy() for k, v in model.state_dict().items()}
    grads = {k: p.grad.numpy() for k, p in model.named_parameters() if p.grad is not None}
    return graphs, bmg, sd, targets.numpy(), mask.numpy(), preds.detach().numpy(), float(loss), grads


def main():
    os.makedirs(OUT_DIR, exist_ok=True)
    for name, (kind, n, seed, cfg) in CASES.items():
        graphs, bmg, sd, targets, mask, preds, loss, grads = run_reference(kind, n, seed, cfg)
        blob = dict(kind=np.asarray(kind), n=np.asarray(n), seed=np.asarray(seed), targets=targets, mask=mask,
                    preds=preds, loss=np.asarray(loss, np.float64),
                    cfg_keys=np.asarray(list(cfg.keys())), cfg_vals=np.asarray([str(v) for v in cfg.values()]),
                    n_atoms=np.asarray(bmg.n_atoms), n_bonds=np.asarray(bmg.n_bonds))
        for k, v in sd.items():
            blob["param:" + k] = v
        for k, v in grads.items():
            blob["grad:" + k] = v
        path = os.path.join(OUT_DIR, name + ".npz")
        np.savez_compressed(path, **blob)
        print(f"{name}: N_a={bmg.n_atoms} N_b={bmg.n_bonds} loss={loss:.6f} -> {os.path.getsize(path) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
