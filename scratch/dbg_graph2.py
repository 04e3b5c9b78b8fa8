import sys, os, traceback
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import torch
from grappa_b200 import ops, _lib, synthetic as S
from grappa_b200.batching import PackedBatch
from grappa_b200.model import GrappaModel
from grappa_b200.train_step import TrainStep
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = GrappaModel.from_default_config()
ts = TrainStep(model, dev, use_cuda_graph=True)
packed = PackedBatch(S.cfg2_batch(int(sys.argv[1]), int(sys.argv[2]), 0))
variant = sys.argv[3]
try:
    if variant == "bench":
        for i in range(4):
            ts.step(packed, load=True); print("step", i, "ok")
    elif variant == "noload":
        ts.step(packed, load=True)
        for i in range(4):
            ts.step(packed, load=False); print("step", i, "ok")
    elif variant == "sync":
        for i in range(4):
            ts.step(packed, load=True); torch.cuda.synchronize(); print("step", i, "ok")
    torch.cuda.synchronize()
    print("DONE", variant)
except Exception as e:
    traceback.print_exc()
    print("last_error:", _lib._raw_lib().grappa_last_error())
