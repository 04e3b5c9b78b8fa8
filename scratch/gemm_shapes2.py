import sys, collections
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import torch
from grappa_b200 import ops, _lib, synthetic as S
from grappa_b200.batching import PackedBatch
from grappa_b200.model import GrappaModel
from grappa_b200.train_step import TrainStep
dev = torch.device("cuda:0")
torch.manual_seed(0)
ts = TrainStep(GrappaModel.from_default_config(), dev, use_cuda_graph=False)
packed = PackedBatch(S.cfg2_batch(32, 50, 0))
for _ in range(3): ts.step(packed)
prof = _lib.Profiler(); _lib.profiler = prof
ts.compute_gradients(packed, load=False); ts.optimizer_step()
torch.cuda.synchronize(); _lib.profiler = None
groups = collections.OrderedDict()
for rec in prof.records:
    if rec[0] != "grappa_gemm": continue
    a = rec[3]
    key = (a[0], a[1], a[2], a[3], a[11], bool(a[10]), bool(a[12]) + bool(a[14]))
    groups.setdefault(key, []).append(rec)
rows = []
for key, recs in groups.items():
    p2 = _lib.Profiler(); p2.records = recs
    ms = p2.replay_ms({"grappa_gemm"})
    rows.append((ms, key, len(recs)))
tot = sum(r[0] for r in rows)
print(f"total gemm ms/step {tot:.3f}")
print(f"{'lay':>3} {'M':>6} {'N':>6} {'K':>6} act bias nres {'n':>3} {'us each':>8} {'ms tot':>7} {'TF':>6}")
for ms, key, n in sorted(rows, key=lambda r: -r[0]):
    fl = 2.0 * key[1] * key[2] * key[3] * n
    print(f"{key[0]:>3} {key[1]:>6} {key[2]:>6} {key[3]:>6} {key[4]:>3} {int(key[5]):>4} {key[6]:>4} {n:>3} {ms/n*1e3:8.1f} {ms:7.3f} {fl/ms/1e9:6.1f}")
