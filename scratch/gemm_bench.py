import sys
sys.path.insert(0, "/root/repo")
import torch
from grappa_b200 import ops
dev = "cuda"
shapes = [(0,1280,512,512),(0,1280,2048,512),(0,1280,512,2048),(0,2560,1536,512),(0,2560,256,256),(0,12160,1536,512),(0,12160,512,512),(0,6080,256,2048),(0,6080,256,256),
          (1,12160,512,1536),(1,12160,512,512),(1,6080,256,256),(1,6080,2048,256),(1,1280,2048,512),
          (2,512,512,12160),(2,1536,512,12160),(2,256,256,6080),(2,256,2048,6080),(2,512,2048,1280),(2,512,512,1280)]
if len(sys.argv) > 1: shapes = [tuple(int(x) for x in a.split(",")) for a in sys.argv[1:]]
def bench(layout, M, N, K, mode, reps=10):
    if layout == 0: A = torch.randn(M, K, device=dev); B = torch.randn(N, K, device=dev)
    elif layout == 1: A = torch.randn(M, K, device=dev); B = torch.randn(K, N, device=dev)
    else: A = torch.randn(K, M, device=dev); B = torch.randn(K, N, device=dev)
    C = torch.empty(M, N, device=dev)
    ops.MATH_MODE = mode
    f = lambda: ops.gemm(layout, M, N, K, A, A.stride(0), B, B.stride(0), C, N)
    for _ in range(3): f()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    with torch.cuda.graph(g, stream=s):
        for _ in range(reps): f()
    g.replay(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(); g.replay(); g.replay(); b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / (2 * reps) * 1e3
print(f"{'lay':>3} {'M':>6} {'N':>6} {'K':>6} {'ffma us':>9} {'tc us':>9} {'ffma TF':>8} {'tc TF':>8}")
for (l, M, N, K) in shapes:
    t0 = bench(l, M, N, K, 0); t1 = bench(l, M, N, K, 1)
    fl = 2.0 * M * N * K
    print(f"{l:>3} {M:>6} {N:>6} {K:>6} {t0:9.1f} {t1:9.1f} {fl/t0/1e6:8.1f} {fl/t1/1e6:8.1f}", flush=True)
