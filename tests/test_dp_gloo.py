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


# ---- the sharded exchange + update of csrc/dp_exchange.cu, its host-side logic (owner ranges, state gather) ----

def test_owner_shards_tile_every_range_and_match_the_library():
    import ctypes
    from grappa_b200 import _lib
    from grappa_b200.dp import shard_of
    L = _lib.lib()
    for lo, hi in ((0, 0), (0, 4), (8, 20), (40, 20836868 // 4 * 4), (569856, 2669568)):
        for world in (1, 2, 3, 4, 8, 16):
            nxt = lo
            for r in range(world):
                first, count = shard_of(lo, hi, r, world)
                assert first == nxt and first % 4 == 0 and count % 4 == 0 and count >= 0
                nxt = first + count
                a, b = ctypes.c_int64(-1), ctypes.c_int64(-1)
                assert L.grappa_dp_shard(lo, hi, r, world, ctypes.byref(a), ctypes.byref(b)) == 0
                assert (a.value, b.value) == (first, count)
            assert nxt == hi
    with pytest.raises(ValueError):
        shard_of(2, 10, 0, 2)
    a, b = ctypes.c_int64(), ctypes.c_int64()
    assert L.grappa_dp_shard(2, 10, 0, 2, ctypes.byref(a), ctypes.byref(b)) != 0


def _adamw_(p, g, m, v, step, lr=1e-3, b1=0.9, b2=0.999, eps=1e-8, wd=0.01, gscale=1.0):
    """The arithmetic of adamw_elem (csrc/dp_exchange.cu) / adamw_kernel on tensors, in place."""
    g = g * gscale
    bc1, bc2s = 1.0 - b1 ** step, (1.0 - b2 ** step) ** 0.5
    pi = p * (1.0 - lr * wd)
    m.mul_(b1).add_((1.0 - b1) * g)
    v.mul_(b2).add_((1.0 - b2) * g * g)
    p.copy_(pi - (lr / bc1) * (m / (v.sqrt() / bc2s + eps)))


def _shard_worker(rank, world, port, out_dir):
    import torch.distributed as dist
    from grappa_b200.dp import gather_shards_, shard_of
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        n = 4 * 61
        ranges = [(0, 40), (40, 44), (44, n)]
        torch.manual_seed(0)
        p_rep = torch.randn(n)
        p_sh = p_rep.clone()
        m_rep, v_rep, m_sh, v_sh = (torch.zeros(n) for _ in range(4))
        for step in range(1, 4):
            torch.manual_seed(100 * step + rank)
            g = torch.randn(n)
            total = g.clone()
            dist.all_reduce(total)                                # what every rank would hold after an all-reduce
            _adamw_(p_rep, total, m_rep, v_rep, step, gscale=1.0 / world)       # replicated: AdamW on everything, on every rank
            for lo, hi in ranges:                                # sharded: the owner updates its elements and sends the parameters
                first, count = shard_of(lo, hi, rank, world)
                sl = slice(first, first + count)
                _adamw_(p_sh[sl], total[sl], m_sh[sl], v_sh[sl], step, gscale=1.0 / world)
                for r in range(world):
                    f, c = shard_of(lo, hi, r, world)
                    if c:
                        dist.broadcast(p_sh[f:f + c], src=r)
            assert torch.equal(p_sh, p_rep), f"step {step}: sharded update differs from the replicated one"
        own = torch.zeros(n, dtype=torch.bool)
        for lo, hi in ranges:
            first, count = shard_of(lo, hi, rank, world)
            own[first:first + count] = True
        assert torch.equal(m_sh[own], m_rep[own]) and not torch.equal(m_sh, m_rep)      # only the owner's part is current
        gather_shards_((m_sh, v_sh), ranges)
        assert torch.equal(m_sh, m_rep) and torch.equal(v_sh, v_rep)
        open(os.path.join(out_dir, f"ok{rank}"), "w").close()
    finally:
        dist.destroy_process_group()


def test_sharded_update_equals_replicated_update_world2(tmp_path):
    port = _free_port()
    mp.spawn(_shard_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert sorted(os.listdir(tmp_path)) == ["ok0", "ok1"]


def test_update_stages_follow_the_forward_order_of_the_default_model():
    """Host logic of the staged exchange (train_step.TrainStep): contiguous stages in the order the forward reads the
    parameters, block 0 cut in front of its self-interaction weights, the four writers last."""
    from grappa_b200.dp import update_stages
    from grappa_b200.model import GrappaModel
    from grappa_b200.train_step import FlatParameters
    model = GrappaModel.from_default_config()
    flat = FlatParameters(model, "cpu")
    n_blocks = len(model.gnn.blocks)
    bounds, marks, writer_stage, split = update_stages(flat.offsets, flat.numel, n_blocks, (0,))
    assert bounds[0] == 0 and bounds[-1] == flat.numel and bounds == sorted(set(bounds)) and all(b % 4 == 0 for b in bounds)
    assert len(bounds) - 1 == 6 and writer_stage == 5 and bounds[5] == split
    assert marks == {(0, 1): 1, (1, 0): 2, (2, 0): 3, (3, 0): 4}
    stage_of = lambda name: max(i for i in range(len(bounds) - 1) if bounds[i] <= flat.offsets[name])
    for name, p in flat.named:                                   # no parameter straddles a boundary
        s = stage_of(name)
        assert flat.offsets[name] + p.numel() <= bounds[s + 1], name
        if name.startswith("gnn.pre_dense"):
            assert s == 0
        elif name.startswith("gnn.att_blocks.0."):
            assert s == (1 if ".self_interaction." in name else 0), name
        elif name.startswith("gnn.att_blocks."):
            assert s == int(name.split(".")[2]) + 1, name
        elif name.startswith("gnn.post_dense"):
            assert s == 4
        else:
            assert s == 5, name
    assert bounds[1] == flat.offsets["gnn.att_blocks.0.self_interaction.0.weight"]
    assert 4 * bounds[1] < 2.5e6 < 10e6 < 4 * bounds[2]          # first stage 2.3 MB instead of 10.7 MB
    # without the extra cut: the five stages of the coarser scheme; every block cut: nine
    assert len(update_stages(flat.offsets, flat.numel, n_blocks, ())[0]) - 1 == 5
    assert len(update_stages(flat.offsets, flat.numel, n_blocks, (0, 1, 2, 3))[0]) - 1 == 9
    # a model without writers has no writer stage
    gnn_only = {n: o for n, o in flat.offsets.items() if n.startswith("gnn.")}
    b2, _, ws2, split2 = update_stages(gnn_only, split, n_blocks, (0,))
    assert ws2 is None and split2 == split and b2[-1] == split


def test_switch_exchange_is_only_offered_for_multi_gpu_cuda_jobs(monkeypatch):
    from grappa_b200.dp import SwitchMemory
    assert SwitchMemory.available(1) is False                     # one rank: nothing to exchange
    monkeypatch.setenv("GRAPPA_DP_SWITCH", "0")
    assert SwitchMemory.available(8) is False                     # explicit opt-out -> NCCL all-reduce path
    monkeypatch.delenv("GRAPPA_DP_SWITCH")
    assert SwitchMemory.available(8) is torch.cuda.is_available() # needs CUDA devices (False on the CPU test box)
