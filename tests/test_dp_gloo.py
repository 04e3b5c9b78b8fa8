"""World-size-2 data-parallel logic on CPU (gloo): the per-rank molecule slices tile the global batch, and summing the
flat gradient buffers over ranks and scaling by 1/world reproduces the gradient of the global molecule-wise loss
(checked with the CPU oracle on a small model-free quadratic and on the real oracle loss)."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from grappa_b200.dp import all_reduce_sum_, flat_views, rank_seeds, rank_slice


def test_rank_slices_tile_the_batch():
    for n, w in ((256, 8), (10, 4), (3, 2), (7, 7)):
        cover = []
        for r in range(w):
            lo, hi = rank_slice(n, r, w)
            cover += list(range(lo, hi))
        assert cover == list(range(n))
    assert rank_seeds(3, 32) == list(range(96, 128))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for p in (root, os.path.join(root, "oracle")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import grappa_oracle as O
    from weights import deterministic_fill_
    from grappa_b200 import synthetic as S
    from grappa_b200.graph import batch
    from grappa_b200.model import GrappaModel
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.distributed.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    sd = GrappaModel.from_default_config().state_dict()
    deterministic_fill_(sd)
    keys = [k for k, v in sd.items() if torch.is_floating_point(v) and not k.startswith("gnn.blocks.") and
            not any(c in k for c in O.OracleTrainer._CONST) and "grappa_transformer.norm." not in k]
    leaf = {k: (v.clone().requires_grad_(True) if k in keys else v.clone()) for k, v in sd.items()}
    n_global = 2 * world
    lo, hi = rank_slice(n_global, rank, world)
    g = batch([S.trimer_molecule(seed=s, n_confs=2) for s in range(lo, hi)])
    loss, _, _ = O.train_step_loss(leaf, g)
    loss.backward()
    flat, views, _ = flat_views([leaf[k] for k in keys])
    for v, k in zip(views, keys):
        v.copy_(leaf[k].grad)
    scale = all_reduce_sum_(flat, world)
    flat.mul_(scale)
    if rank == 0:
        # single-process reference: the global batch in one go
        leaf2 = {k: (v.clone().requires_grad_(True) if k in keys else v.clone()) for k, v in sd.items()}
        gg = batch([S.trimer_molecule(seed=s, n_confs=2) for s in range(n_global)])
        loss2, _, _ = O.train_step_loss(leaf2, gg)
        loss2.backward()
        worst = 0.0
        for v, k in zip(views, keys):
            r = leaf2[k].grad
            worst = max(worst, float((v - r).abs().max() / (r.abs().max() + 1e-12)))
        np.save(os.path.join(out_dir, "worst.npy"), np.array([worst, float(loss2)]))
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()


def test_two_rank_gradient_allreduce_matches_global_batch(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    worst, loss = np.load(os.path.join(str(tmp_path), "worst.npy"))
    assert np.isfinite(loss)
    assert worst < 2e-4, f"data-parallel mean gradient deviates from the global-batch gradient by {worst:.2e} (relative to max)"
