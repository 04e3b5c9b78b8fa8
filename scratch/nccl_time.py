"""All-reduce device time at the bucket sizes of the staged update (development probe; run under torchrun)."""
import os, torch, torch.distributed as dist
rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local); dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
for n in (660_000, 2_640_000, 10_680_000, 20_840_000):
    x = torch.randn(n, device=dev)
    for _ in range(5): dist.all_reduce(x)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20): dist.all_reduce(x)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 20
    if rank == 0:
        print(f"world {world}: all_reduce {n * 4 / 1e6:.1f} MB: {ms * 1e3:.1f} us  busbw {2 * (world - 1) / world * n * 4 / ms / 1e6:.0f} GB/s", flush=True)
dist.barrier(); dist.destroy_process_group()
