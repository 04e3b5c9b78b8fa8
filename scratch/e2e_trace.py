"""Host-side timing of the store-fed step at world > 1 (development probe): torchrun ... scratch/e2e_trace.py"""
import os, sys, time
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import numpy as np, torch
from grappa_b200 import ops, synthetic as S
from grappa_b200.batching import PackedBatch
from grappa_b200.model import GrappaModel
from grappa_b200.store import MoleculeStore
from grappa_b200.train_step import TrainStep
world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local); dev = torch.device("cuda", local)
if world > 1:
    torch.distributed.init_process_group("nccl", device_id=dev)
torch.manual_seed(0)
ts = TrainStep(GrappaModel.from_default_config(), dev, world_size=world)
store = MoleculeStore(S.cfg2_graphs(96, 50, first_seed=1000 + 96 * rank), dev)
rng = np.random.default_rng(rank)
draws = [rng.choice(store.n, 32, replace=False) for _ in range(30)]
for ids in draws[:8]:
    ts.step_from_store(store, ids)
torch.cuda.synchronize()
T = {k: 0.0 for k in ("prepare", "state", "collate", "grads", "update", "loss")}
t_all = time.perf_counter()
pending = None
for ids in draws[8:]:
    t0 = time.perf_counter(); spec = store.prepare(ids); t1 = time.perf_counter()
    st = ts._state_for(spec); t2 = time.perf_counter()
    store.collate_into(spec, st["batch"]); t3 = time.perf_counter()
    loss = ts.compute_gradients(spec, load=False); t4 = time.perf_counter()
    ts.exchange_and_update(); t5 = time.perf_counter()
    h = ts.read_loss_async(loss)
    if pending is not None: pending.value()
    pending = h; t6 = time.perf_counter()
    for k, v in zip(T, (t1 - t0, t2 - t1, t3 - t2, t4 - t3, t5 - t4, t6 - t5)): T[k] += v
pending.value(); torch.cuda.synchronize()
n = len(draws) - 8
print(f"rank {rank}: total {(time.perf_counter() - t_all) / n * 1e3:.2f} ms/step; " + ", ".join(f"{k} {v / n * 1e3:.2f}" for k, v in T.items()), flush=True)
if world > 1:
    torch.distributed.barrier(); torch.distributed.destroy_process_group()
