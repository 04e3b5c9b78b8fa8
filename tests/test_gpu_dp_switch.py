"""The fused NVSwitch exchange + AdamW kernel (grappa_dp_exchange_adamw, csrc/dp_exchange.cu) against the NCCL
all-reduce + grappa_adamw path it replaces, on 2 GPUs (skipped on a one-GPU box): same weights, same per-rank batches,
dropout off, a few optimisation steps.  world = 2 makes the gradient sum order-free (a + b), so the two paths may differ
only by the rounding of the AdamW arithmetic itself."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu


def _worker(rank, world, port, ret):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world),
                      GRAPPA_DP_SPIN_SECONDS="20")
    import torch.distributed as dist
    from grappa_b200 import synthetic as S
    from grappa_b200.batching import PackedBatch
    from grappa_b200.dp import shard_of
    from grappa_b200.graph import batch
    from grappa_b200.model import GrappaModel
    from grappa_b200.train_step import TrainStep
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    dist.init_process_group("nccl", device_id=dev)
    try:
        packed = PackedBatch(batch(S.cfg2_graphs(4, 5, first_seed=10 * rank)), pin=False)
        results = {}
        for name, switch, split in (("nccl", False, True), ("switch", True, True), ("switch_serial", True, False)):
            torch.manual_seed(0)
            model = GrappaModel.from_default_config()
            for m in model.modules():
                if isinstance(m, torch.nn.Dropout):
                    m.p = 0.0
            ts = TrainStep(model, dev, lr=1e-3, weight_decay=0.01, world_size=world, use_cuda_graph=False, split_update=split,
                           switch_exchange=switch)
            assert (ts._switch is not None) == switch
            losses = [float(ts.step(packed)) for _ in range(4)]
            ts.synchronize()
            ts.gather_optimizer_state()
            results[name] = (losses, ts.flat.params.clone(), ts.flat.exp_avg.clone(), ts.flat.exp_avg_sq.clone())
            # every rank holds the same parameters, bit for bit
            mine = ts.flat.params.view(torch.int32).to(torch.int64).sum().reshape(1)
            both = [torch.zeros_like(mine) for _ in range(world)]
            dist.all_gather(both, mine)
            assert all(torch.equal(b, both[0]) for b in both), f"{name}: ranks hold different parameters"
            del ts
        ref = results["nccl"]
        for name in ("switch", "switch_serial"):
            got = results[name]
            for a, b in zip(ref[0], got[0]):
                assert abs(a - b) <= 1e-5 * abs(a), (name, ref[0], got[0])
            for what, a, b in (("params", ref[1], got[1]), ("exp_avg", ref[2], got[2]), ("exp_avg_sq", ref[3], got[3])):
                err = float((a - b).abs().max() / a.abs().max())
                assert err < 2e-5, f"{name}: {what} differ from the NCCL path by {err:.2e}"
        # the parameters did move (lr 1e-3, 4 steps)
        torch.manual_seed(0)
        p0 = torch.cat([p.detach().reshape(-1) for p in GrappaModel.from_default_config().parameters()])
        assert float(ref[1].abs().sum()) != float(p0.abs().sum())
        # host-side shard arithmetic matches the library's
        assert shard_of(0, 40, 1, 2) == (20, 20) and shard_of(8, 20, 0, 2) == (8, 4)
        ret[rank] = "ok"
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs 2 GPUs on one NVSwitch domain")
def test_switch_exchange_matches_nccl_path():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 29600 + os.getpid() % 300
    procs = [ctx.Process(target=_worker, args=(r, 2, port, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=600)
    for p in procs:
        if p.is_alive():
            p.terminate()
    assert [p.exitcode for p in procs] == [0, 0], f"worker exit codes {[p.exitcode for p in procs]}"
    assert dict(ret) == {0: "ok", 1: "ok"}
