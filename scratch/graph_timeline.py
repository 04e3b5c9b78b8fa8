import sys
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/oracle")
import torch
from grappa_b200 import ops, synthetic as S, model as M
from grappa_b200.batching import PackedBatch
from grappa_b200.model import GrappaModel
from grappa_b200.train_step import TrainStep
dev = torch.device("cuda:0")
torch.manual_seed(0)
ts = TrainStep(GrappaModel.from_default_config(), dev, use_cuda_graph=True)
packed = PackedBatch(S.cfg2_batch(32, 50, 0))
MARKS = []
ACTIVE = [False]
def mark(name):
    if ACTIVE[0]:
        ev = torch.cuda.Event(enable_timing=True, external=True)
        ev.record()
        MARKS.append((name, ev))
gnn_fwd = ts.model.gnn.forward
def gnn_wrapped(g, plan=None):
    mark("start")
    out = gnn_fwd(g, plan)
    mark("gnn fwd done")
    h = out.nodes["n1"].data["h"]
    if h.requires_grad:
        h.register_hook(lambda gr: (mark("dL/dh ready (writers bwd done on main)"), gr)[1])
    return out
ts.model.gnn.forward = gnn_wrapped
pw_fwd = ts.model.parameter_writer.forward
def pw_wrapped(g, plan):
    out = pw_fwd(g, plan)
    mark("writers fwd done")
    return out
ts.model.parameter_writer.forward = pw_wrapped
# per-writer marks on their own streams
for name in ("bond_writer", "angle_writer", "proper_writer", "improper_writer"):
    w = getattr(ts.model.parameter_writer, name)
    f = w.forward
    def make(f, name):
        def wrapped(g, plan):
            mark(f"  {name} fwd start")
            out = f(g, plan)
            mark(f"  {name} fwd end")
            return out
        return wrapped
    w.forward = make(f, name)
en_fwd = ts.energy.forward
def en_wrapped(g):
    out = en_fwd(g)
    mark("energy fwd done")
    return out
ts.energy.forward = en_wrapped
orig_join = ops.join_side_work
def join_wrapped(device):
    mark("main chain bwd done (before side join)")
    orig_join(device)
    mark("side streams joined")
ops.join_side_work = join_wrapped
for _ in range(3): ts.step(packed)
ACTIVE[0] = True
ts.step(packed)          # capture (marks recorded as external event nodes)
ACTIVE[0] = False
for _ in range(3): ts.step(packed)
torch.cuda.synchronize()
t0 = MARKS[0][1]
for name, ev in MARKS:
    print(f"{t0.elapsed_time(ev)*1e3:9.1f} us  {name}")
