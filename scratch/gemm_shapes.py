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
res = {}
for mode in (0, 1):
    ops.MATH_MODE = mode
    for _ in range(3): ts.step(packed)
    _lib.profiler = _lib.Profiler()
    for _ in range(3): ts.compute_gradients(packed, load=False)
    torch.cuda.synchronize()
    agg = collections.OrderedDict()
    for name, a, b, args in _lib.profiler.records:
        if name != "grappa_gemm": continue
        key = (args[0], args[1], args[2], args[3])
        d = agg.setdefault(key, [0, 0.0]); d[0] += 1; d[1] += a.elapsed_time(b)
    _lib.profiler = None
    res[mode] = agg
print(f"{'layout':>6} {'M':>6} {'N':>6} {'K':>6} {'n':>3} {'ffma us':>9} {'tc us':>9} {'ffma TF':>8} {'tc TF':>8}")
tot = [0, 0]
for key in res[0]:
    n, t0 = res[0][key]; _, t1 = res[1][key]
    fl = 2.0 * key[1] * key[2] * key[3]
    tot[0] += t0; tot[1] += t1
    print(f"{key[0]:>6} {key[1]:>6} {key[2]:>6} {key[3]:>6} {n//3:>3} {t0/n*1e3:9.1f} {t1/n*1e3:9.1f} {fl/(t0/n*1e-3)/1e12:8.1f} {fl/(t1/n*1e-3)/1e12:8.1f}")
print("total ms per step", tot[0]/3, tot[1]/3)
