"""Where does the multi-GPU step lose time?  torchrun --nproc-per-node N scratch/dp_probe.py  (development probe)
Variants of the gradient exchange on the cfg-2 step, each timed like bench.py (device-resident inputs, CUDA graph):
  staged      the default: five forward-ordered stages, stages 1..4 under the next forward
  nocomm      the same choreography with the collectives left out (results meaningless: cost of the structure alone)
  serial      one all-reduce + one AdamW after backward (split_update off)
  serial_nocomm  that, without the collective (= the one-GPU step)
"""
import os, sys
sys.path.insert(0, "/root/repo")
import torch
from grappa_b200 import ops
from grappa_b200.batching import PackedBatch
from grappa_b200.model import GrappaModel
from grappa_b200.train_step import TrainStep
import grappa_b200.train_step as TS
import bench

world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
if world > 1:
    torch.distributed.init_process_group("nccl", device_id=dev)
dist = torch.distributed
real_all_reduce = dist.all_reduce
K = int(os.environ.get("K", "40"))


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def run(name, split, comm, ffn=(0,), switch=False):
    torch.manual_seed(0)
    model = GrappaModel.from_default_config()
    ts = TrainStep(model, dev, lr=1e-5, world_size=world, split_update=split, ffn_split_blocks=ffn, switch_exchange=switch)
    ops.DropoutSeeds.manual_seed(1234 + rank)
    packed = PackedBatch(bench.build_batch(first_seed=32 * rank), pin=True)
    dist.all_reduce = real_all_reduce if comm else (lambda *a, **k: None)
    if not comm:
        TS.all_reduce_sum_ = lambda flat, w, pg: 1.0 / w
    try:
        for _ in range(8):
            ts.step(packed)
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = []
        for rep in range(3):
            barrier()
            a.record()
            for _ in range(K):
                ts.step(packed, load=False)
            b.record()
            barrier()
            t = torch.tensor([a.elapsed_time(b) / K], device=dev)
            if world > 1:
                real_all_reduce(t, op=dist.ReduceOp.MAX)
            best.append(float(t))
        ts.synchronize()
        if rank == 0:
            print(f"{name:16s} N={world}  ms/step (max over ranks, 3 repeats): " + "  ".join(f"{x:.4f}" for x in best), flush=True)
    finally:
        dist.all_reduce = real_all_reduce
    del ts, model
    torch.cuda.empty_cache()


orig_ars = TS.all_reduce_sum_
ALL = (("switch6", True, True, (0,), True), ("switch5", True, True, (), True), ("switch_serial", False, True, (), True),
       ("staged5", True, True, ()), ("staged6", True, True, (0,)), ("staged9", True, True, (0, 1, 2, 3)),
       ("nocomm6", True, False, (0,)), ("serial", False, True, ()), ("serial_nocomm", False, False, ()))
only = os.environ.get("VARIANTS", "").split(",") if os.environ.get("VARIANTS") else None
for name, split, comm, ffn, *sw in ALL:
    if only is not None and name not in only:
        continue
    TS.all_reduce_sum_ = orig_ars
    run(name, split, comm, ffn, bool(sw and sw[0]))
if world > 1:
    dist.destroy_process_group()
