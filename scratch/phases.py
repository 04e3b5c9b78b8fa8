import sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import torch
from grappa_b200 import ops, _lib, synthetic as S, model as M
from grappa_b200.batching import PackedBatch
from grappa_b200.model import GrappaModel
from grappa_b200.train_step import TrainStep
from grappa_b200.plan import get_plan
dev = torch.device("cuda:0")
torch.manual_seed(0)
ts = TrainStep(GrappaModel.from_default_config(), dev, use_cuda_graph=False)
packed = PackedBatch(S.cfg2_batch(32, 50, 0))
for _ in range(3): ts.step(packed)
g = ts._state_for(packed)["batch"].graph
model, en = ts.model, ts.energy
plan = get_plan(g)

def timed(name, fn, grad=True):
    side = torch.cuda.Stream()
    ops.GRAD_SINK = ts.flat.sink; ops.DropoutSeeds.dev = ts.seed_dev
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(2): fn()
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph, stream=side):
        fn()
    graph.replay(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): graph.replay()
    b.record(); torch.cuda.synchronize()
    ops.GRAD_SINK = None; ops.DropoutSeeds.dev = None
    print(f"{name:40s} {a.elapsed_time(b)/10:8.3f} ms")

def gnn_fwd():
    with torch.no_grad(): model.gnn(g, plan)
def model_fwd():
    with torch.no_grad(): model(g)
def model_fwd_serial():
    M.PARALLEL_WRITERS = False
    with torch.no_grad(): model(g)
    M.PARALLEL_WRITERS = True
def writer_only(w):
    def f():
        with torch.no_grad(): w(g, plan)
    return f
def fwd_energy():
    with torch.no_grad(): en(model(g))
def gnn_fwd_bwd():
    h = model.gnn(g, plan).nodes["n1"].data["h"]
    h.backward(torch.ones_like(h)); ops.join_side_work(dev)
def full():
    ts._fwd_bwd(g)
def full_serial():
    M.PARALLEL_WRITERS = False
    ts._fwd_bwd(g)
    M.PARALLEL_WRITERS = True
timed("gnn forward", gnn_fwd)
timed("model forward (parallel writers)", model_fwd)
timed("model forward (serial writers)", model_fwd_serial)
model(g)
for name, w in (("bond", model.parameter_writer.bond_writer), ("angle", model.parameter_writer.angle_writer), ("proper", model.parameter_writer.proper_writer), ("improper", model.parameter_writer.improper_writer)):
    timed(f"  {name} writer forward alone", writer_only(w))
timed("model + energy forward", fwd_energy)
timed("gnn forward+backward", gnn_fwd_bwd)
import sys as _s; _s.stdout.flush()
try:
    timed("full fwd+loss+bwd (parallel)", full)
except Exception as e: print("full parallel failed:", str(e)[:80])
try:
    timed("full fwd+loss+bwd (serial writers)", full_serial)
except Exception as e: print("full serial failed:", str(e)[:80])
hleaf = g.nodes["n1"].data["h"].detach().clone().requires_grad_(True)
def writer_fb(w, lvl):
    def f():
        g.nodes["n1"].data["h"] = hleaf
        w(g, plan)
        d = g.nodes[lvl].data
        outs = [d[k] for k in ("k", "eq") if k in d and d[k].requires_grad]
        torch.autograd.backward(outs, [torch.ones_like(o) for o in outs])
        ops.join_side_work(dev)
    return f
for name, w, lvl in (("bond", model.parameter_writer.bond_writer, "n2"), ("angle", model.parameter_writer.angle_writer, "n3"), ("proper", model.parameter_writer.proper_writer, "n4"), ("improper", model.parameter_writer.improper_writer, "n4_improper")):
    try:
        timed(f"  {name} writer fwd+bwd alone", writer_fb(w, lvl))
    except Exception as e:
        print(name, "failed", str(e)[:100])
