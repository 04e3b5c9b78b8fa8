"""One launch of every distinct GEMM (entry point, shape, class) of a cfg-2 training step, in a fixed order, between
cudaProfilerStart/Stop -- to be run under `ncu --profile-from-start off --set full -k regex:gemm_`; writes the order
(one line per launch that ncu sees, split-K reduce kernels included as their own lines) to argv[1]."""
import sys, ctypes, collections
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
first = collections.OrderedDict()
for n, _, _, a in prof.records:
    if n in names and key(n, a) not in first:
        first[key(n, a)] = (n, a)
raw = _lib._raw_lib()
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
torch.cuda.profiler.start()
with open(sys.argv[1], "w") as f:
    f.write("entry,layout,M,N,K,class\n")
    for k, (n, a) in first.items():
        rc = getattr(raw, n)(*a[:-1], st)
        assert rc == 0
        f.write(",".join(str(x) for x in k) + "\n")
torch.cuda.synchronize()
torch.cuda.profiler.stop()
