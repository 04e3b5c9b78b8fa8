"""Per-shape table of the GEMMs of one cfg-2 training step: python scratch/gemm_table.py [out.csv]
entry point, layout, M, N, K, class (fg = critical chains, bg = side-stream weight gradients), calls, tiles, us per call
(those launches replayed back to back in a CUDA graph), TFLOP/s, algorithmic MB per call (fp32 A + B + C once)."""
import sys, collections
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import torch
from grappa_b200 import _lib, ops, synthetic as S
from grappa_b200.batching import PackedBatch
from grappa_b200.model import GrappaModel
from grappa_b200.train_step import TrainStep
dev = torch.device("cuda:0")
torch.manual_seed(0)
ts = TrainStep(GrappaModel.from_default_config(), dev, use_cuda_graph=False)
packed = PackedBatch(S.cfg2_batch(32, 50), pin=True)
for _ in range(2):
    ts.step(packed)
prof = _lib.Profiler(); _lib.profiler = prof
ts.compute_gradients(packed, load=False); ts.optimizer_step()
torch.cuda.synchronize(); _lib.profiler = None
names = {"grappa_gemm", "grappa_gemm_ex", "grappa_gemm_actgrad", "grappa_gemm_bf16x3"}
def key(n, a):
    if n == "grappa_gemm_bf16x3":
        return (n, "NT(planes)", a[0], a[1], a[2], "bg" if (a[19] >> 16) & 1 else "fg")
    mm = a[19] if n != "grappa_gemm_actgrad" else a[12]
    return (n, ("NT", "NN", "TN")[a[0]], a[1], a[2], a[3], "bg" if (mm >> 16) & 1 else "fg")
groups = collections.OrderedDict()
for n, _, _, a in prof.records:
    if n in names:
        groups.setdefault(key(n, a), 0)
        groups[key(n, a)] += 1
rows = []
for k, cnt in groups.items():
    ms = prof.replay_ms(names, keep=lambda n, a, k=k: key(n, a) == k)
    n, lay, M, N, K, cls = k
    fl = 2.0 * M * N * K
    us = ms * 1e3 / cnt
    tiles = ((M + 127) // 128) * ((N + 127) // 128)
    rows.append((n, lay, M, N, K, cls, cnt, tiles, us, fl / us / 1e6, 4.0 * (M * K + N * K + M * N) / 1e6))
rows.sort(key=lambda r: -r[8] * r[6])
out = open(sys.argv[1], "w") if len(sys.argv) > 1 else sys.stdout
print("entry,layout,M,N,K,class,calls,tiles_128x128,us_per_call,TFLOP_per_s,algorithmic_MB_per_call,total_us", file=out)
for r in rows:
    print(",".join(str(x) if not isinstance(x, float) else f"{x:.2f}" for x in r) + f",{r[8] * r[6]:.1f}", file=out)
tot = sum(r[8] * r[6] for r in rows); fl = sum(2.0 * r[2] * r[3] * r[4] * r[6] for r in rows)
print(f"# total {tot / 1e3:.3f} ms serial, {fl / 1e9:.1f} GFLOP, {fl / tot / 1e6:.1f} TFLOP/s over all launches", file=out)
