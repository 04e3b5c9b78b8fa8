"""Does an NCCL all-reduce captured inside a CUDA graph replay correctly here? (development probe, run under torchrun)"""
import os, sys, time
import torch, torch.distributed as dist
rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
mode = sys.argv[1] if len(sys.argv) > 1 else "thread_local"
x = torch.ones(10_000_000, device=dev) * (rank + 1)
y = torch.zeros_like(x)
side = torch.cuda.Stream(device=dev); comm = torch.cuda.Stream(device=dev)
for _ in range(3):                       # eager warm-up (communicator creation outside capture)
    dist.all_reduce(x.clone())
torch.cuda.synchronize()
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g, stream=side, capture_error_mode=mode):
    y.copy_(x)
    y.mul_(2.0)
    comm.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(comm):        # collective on a forked stream, as the bucketed exchange would issue it
        dist.all_reduce(y)
    z = x * 3.0                          # independent work that may overlap
    torch.cuda.current_stream().wait_stream(comm)
    out = y + z
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(20):
    g.replay()
torch.cuda.synchronize()
dt = (time.perf_counter() - t0) / 20
expect = 2.0 * sum(r + 1 for r in range(world)) + 3.0 * (rank + 1)
print(f"rank {rank} mode {mode}: out[0] = {float(out[0])} expect {expect} ok={float(out[0]) == expect} {dt * 1e3:.3f} ms/replay", flush=True)
dist.barrier(); dist.destroy_process_group()
