import sys, time
sys.path.insert(0, "/root/repo")
import torch
from grappa_b200 import _lib, ops, synthetic as S
from grappa_b200.model import GrappaModel
dev = torch.device("cuda:0")
torch.manual_seed(0)
model = GrappaModel.from_default_config().to(dev).eval()
g = S.cfg3_graph().to(dev)
with torch.no_grad():
    for _ in range(3): model(g)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10): model(g)
    t_host = (time.perf_counter() - t0) / 10
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): model(g)
    b.record(); torch.cuda.synchronize()
    print(f"wall {a.elapsed_time(b)/10:.2f} ms, host enqueue time {t_host*1e3:.2f} ms")
    prof = _lib.Profiler(); _lib.profiler = prof
    model(g)
    summ = prof.summary(); _lib.profiler = None
    tot = 0
    for n in sorted(summ):
        ms = prof.replay_ms({n}); tot += ms
        print(f"{n:32s} {summ[n]['calls']:4d} calls {ms:8.3f} ms" + (f"  {summ[n]['flops']/ms/1e9:7.1f} TF" if summ[n]['flops'] else ""))
    print("sum", tot)
    # whole forward as one CUDA graph
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        model(g)
    torch.cuda.synchronize()
    gr = torch.cuda.CUDAGraph()
    with torch.cuda.graph(gr, stream=side):
        model(g)
    gr.replay(); torch.cuda.synchronize()
    a.record()
    for _ in range(10): gr.replay()
    b.record(); torch.cuda.synchronize()
    print(f"graph replay {a.elapsed_time(b)/10:.2f} ms")
