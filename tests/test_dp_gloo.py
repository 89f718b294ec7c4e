"""CPU, world_size 2, gloo: the host-side logic of the data-parallel path --
molecule sharding and the single flat-buffer gradient all-reduce.  (The encoder kernels themselves
need a GPU; here the per-rank gradients come from the CPU oracle on each rank's shard, so the test
also checks the property the DP path relies on: gradients of disjoint molecule shards add up.)"""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from polymer_chemprop_b200 import BatchMolGraph, dp, synth


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _weights(h, fa, fb):
    g = torch.Generator().manual_seed(0)
    return {"W_i.weight": torch.randn(h, fb, generator=g) * 0.1, "W_h.weight": torch.randn(h, h, generator=g) * 0.1,
            "W_o.weight": torch.randn(h, fa + h, generator=g) * 0.1, "W_o.bias": torch.randn(h, generator=g) * 0.1}


def _oracle_grads(graphs, d_out, h):
    from oracle import mpn_oracle
    bmg = BatchMolGraph(graphs)
    w = _weights(h, bmg.atom_fdim, bmg.bond_fdim)
    _out, grads = mpn_oracle.encoder_forward_backward(bmg.get_components(), w, d_out, depth=3)
    return w, grads


def _worker(rank, world, port, n_graphs, h, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.set_num_threads(2)
        graphs = synth.make_graphs("copolymer", n_graphs, seed=7, dense_features=True)
        d_out = torch.randn(n_graphs, h, generator=torch.Generator().manual_seed(1))
        idx = dp.shard_range(n_graphs, rank, world)
        w, grads = _oracle_grads([graphs[i] for i in idx], d_out[idx.start:idx.stop], h)
        params = [torch.nn.Parameter(v.clone()) for v in w.values()]
        for p, k in zip(params, w):
            p.grad = grads[k].clone()
        reducer = dp.FlatGradAllReduce(params, average=False)
        flat = reducer()
        assert flat.numel() == sum(p.numel() for p in params)
        torch.save({k: p.grad for k, p in zip(w, params)}, os.path.join(out_dir, f"rank{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_shard_range_partitions_equally():
    for n, world in [(4096, 8), (10, 3), (7, 2)]:
        shards = [dp.shard_range(n, r, world) for r in range(world)]
        assert all(len(s) == n // world for s in shards)
        seen = [i for s in shards for i in s]
        assert seen == list(range((n // world) * world))
    with pytest.raises(ValueError):
        dp.shard_range(4, 2, 2)


def test_flat_gradient_allreduce_world2_gloo(tmp_path):
    world, n_graphs, h = 2, 8, 24
    port = _free_port()
    mp.spawn(_worker, args=(world, port, n_graphs, h, str(tmp_path)), nprocs=world, join=True)
    r0, r1 = (torch.load(tmp_path / f"rank{r}.pt") for r in range(world))
    graphs = synth.make_graphs("copolymer", n_graphs, seed=7, dense_features=True)
    d_out = torch.randn(n_graphs, h, generator=torch.Generator().manual_seed(1))
    _w, full = _oracle_grads(graphs, d_out, h)          # single-process gradient of the whole batch
    for k in full:
        assert torch.equal(r0[k], r1[k]), k             # every rank ends with the same reduced gradient
        err = float((r0[k] - full[k]).norm() / full[k].norm())
        assert err <= 1e-5, (k, err)                    # == sum over disjoint molecule shards
