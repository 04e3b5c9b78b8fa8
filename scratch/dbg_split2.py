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
order = [(a[0] == "g", bool(int(a[1]))) for a in sys.argv[1:]]
for use_graph, overlap in order:
    torch.manual_seed(0)
    ops.DropoutSeeds.manual_seed(1234)
    ts = TrainStep(GrappaModel.from_default_config(), dev, use_cuda_graph=use_graph, overlap_comm=overlap)
    for i in range(5):
        try:
            l = float(ts.step(packed))
            print("graph", use_graph, "overlap", overlap, "step", i, "ok", l)
        except Exception as e:
            print("graph", use_graph, "overlap", overlap, "step", i, "FAILED", str(e)[:100]); sys.exit(0)
