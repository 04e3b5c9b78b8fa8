import sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle"); sys.path.insert(0, "/root/repo/tests")
import torch
import test_gpu_model as T
from grappa_b200 import ops, synthetic as S
from grappa_b200.batching import PackedBatch
from grappa_b200.graph import batch
from grappa_b200.train_step import TrainStep
packed = PackedBatch(batch([S.trimer_molecule(seed=s, n_confs=5) for s in range(4)]))
variant = sys.argv[1]
for use_graph in ((False, True) if variant != "gonly" else (True,)):
    for overlap in (False, True):
        model, _ = T.build_model(train=True) if variant != "rand" else (T.GrappaModel.from_default_config(), None)
        ops.DropoutSeeds.manual_seed(1234)
        ts = TrainStep(model, T.DEV, use_cuda_graph=use_graph, overlap_comm=overlap)
        for i in range(5):
            try:
                l = float(ts.step(packed))
            except Exception as e:
                print("graph", use_graph, "overlap", overlap, "step", i, "FAILED", str(e)[:100]); sys.exit(0)
        print("graph", use_graph, "overlap", overlap, "ok", l)
