import sys, os
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import torch
from grappa_b200 import ops, _lib, synthetic as S
from grappa_b200.batching import PackedBatch, DeviceBatch
from grappa_b200.model import GrappaModel
from grappa_b200.energy import Energy
from grappa_b200.train_step import TrainStep
from grappa_b200.plan import get_plan

dev = torch.device("cuda:0")
torch.manual_seed(0)
model = GrappaModel.from_default_config()
ts = TrainStep(model, dev, use_cuda_graph=False)
packed = PackedBatch(S.cfg2_batch(int(sys.argv[1]), int(sys.argv[2]), 0))
for _ in range(2):
    ts.step(packed)
torch.cuda.synchronize()
g = ts._state_for(packed)["batch"].graph
en = ts.energy

def try_capture(name, fn, mode="global"):
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    try:
        with torch.cuda.graph(graph, capture_error_mode=mode):
            fn()
        graph.replay(); torch.cuda.synchronize()
        print("OK  ", name, mode)
    except Exception as e:
        print("FAIL", name, mode, str(e).split("\n")[0][:150])
        torch.cuda.synchronize()

def fwd_nograd():
    with torch.no_grad():
        model(g)
def fwd_en_nograd():
    with torch.no_grad():
        en(model(g))
def fwd_grad():
    model(g)
def gnn_bwd():
    ops.GRAD_SINK = ts.flat.sink
    out = model.gnn(g).nodes["n1"].data["h"]
    out.sum().backward()
    ops.GRAD_SINK = None
def lin_bwd():
    x = torch.randn(64, 512, device=dev, requires_grad=True)
    w = torch.randn(256, 512, device=dev, requires_grad=True)
    ops.linear(x, w).sum().backward()
def full():
    ts._fwd_bwd(g)
for mode in ("global", "thread_local"):
    for name, fn in (("fwd_nograd", fwd_nograd), ("fwd_en_nograd", fwd_en_nograd), ("fwd_grad", fwd_grad), ("lin_bwd", lin_bwd), ("gnn_bwd", gnn_bwd), ("full", full)):
        try_capture(name, fn, mode)
