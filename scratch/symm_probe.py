"""Is NVSwitch multicast (symmetric memory) usable on this box, and how fast is an in-switch all-reduce?  (development probe)
torchrun --nproc-per-node N scratch/symm_probe.py"""
import os, sys, time
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as sm

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
n = 20_836_868
t = sm.empty(n, dtype=torch.float32, device=dev)
h = sm.rendezvous(t, dist.group.WORLD)
if rank == 0:
    print("multicast_ptr:", hex(h.multicast_ptr), "buffers:", [hex(p) for p in h.buffer_ptrs], "signal pad bytes:", h.signal_pad_size, flush=True)
gname = dist.group.WORLD.group_name


def timed(name, fn, reps=20):
    for _ in range(3):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record(); torch.cuda.synchronize()
    ms = torch.tensor([a.elapsed_time(b) / reps], device=dev)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"{name:44s} {float(ms)*1e3:9.1f} us", flush=True)


t.fill_(rank + 1.0)
torch.cuda.synchronize(); dist.barrier()
try:
    torch.ops.symm_mem.multimem_all_reduce_(t, "sum", gname)
    torch.cuda.synchronize()
    if rank == 0:
        print("multimem_all_reduce_ result", float(t[0]), float(t[-1]), "expected", world * (world + 1) / 2, flush=True)
except Exception as e:
    print("multimem_all_reduce_ failed:", str(e)[:300], flush=True)
plain = torch.zeros(n, device=dev)
for size in (43_008, 570_000, 2_700_000, n):
    v, p = t[:size], plain[:size]
    timed(f"nccl all_reduce {size*4/1e6:.2f} MB", lambda: dist.all_reduce(p))
    try:
        timed(f"multimem_all_reduce_ {size*4/1e6:.2f} MB", lambda: torch.ops.symm_mem.multimem_all_reduce_(v, "sum", gname))
    except Exception as e:
        print("multimem failed", str(e)[:200])
    try:
        timed(f"two_shot_all_reduce_ {size*4/1e6:.2f} MB", lambda: torch.ops.symm_mem.two_shot_all_reduce_(v, "sum", gname))
    except Exception as e:
        print("two_shot failed", str(e)[:200])
dist.destroy_process_group()
