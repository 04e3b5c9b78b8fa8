import sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import torch
from grappa_b200 import ops, synthetic as S
from grappa_b200.batching import PackedBatch
from grappa_b200.graph import batch
from grappa_b200.model import GrappaModel
from grappa_b200.train_step import TrainStep
dev = torch.device("cuda:0")
packed = PackedBatch(batch([S.trimer_molecule(seed=s, n_confs=5) for s in range(4)]))
ts = TrainStep(GrappaModel.from_default_config(), dev, use_cuda_graph=True, overlap_comm=True)
st = ts._state_for(packed)
g = st["batch"].graph
for i in range(3):
    ts.step(packed)
torch.cuda.synchronize()
print("warm ok")
ga, gb = torch.cuda.CUDAGraph(), torch.cuda.CUDAGraph()
try:
    with torch.cuda.graph(ga, stream=ts._side):
        loss, h, leaf = ts._phase_a(g)
    print("capture A ok")
except Exception as e:
    print("capture A failed:", str(e)[:120]); sys.exit(0)
try:
    with torch.cuda.graph(gb, stream=ts._side, pool=ga.pool()):
        ts._phase_b(h, leaf)
    print("capture B ok")
except Exception as e:
    print("capture B failed:", str(e)[:120])
