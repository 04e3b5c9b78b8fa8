import sys, math
sys.path.insert(0, "/root/repo")
import torch
from grappa_b200 import ops
dev = "cuda"
torch.manual_seed(0)

def run(layout, M, N, K, mode, act=0, bias=False, res=False):
    g = torch.Generator().manual_seed(M + 3 * N + 7 * K + layout)
    if layout == 0:
        A = torch.randn(M, K, generator=g); B = torch.randn(N, K, generator=g); ref = A.double() @ B.double().t()
    elif layout == 1:
        A = torch.randn(M, K, generator=g); B = torch.randn(K, N, generator=g); ref = A.double() @ B.double()
    else:
        A = torch.randn(K, M, generator=g); B = torch.randn(K, N, generator=g); ref = A.double().t() @ B.double()
    bs = torch.randn(N, generator=g) if bias else None
    rs = torch.randn(M, N, generator=g) if res else None
    if bias: ref = ref + bs.double()
    if act: ref = torch.nn.functional.elu(ref)
    if res: ref = ref + rs.double()
    Ad, Bd = A.to(dev), B.to(dev)
    C = torch.full((M, N), float("nan"), device=dev)
    ops.MATH_MODE = mode
    ops.gemm(layout, M, N, K, Ad, Ad.stride(0), Bd, Bd.stride(0), C, N, bs.to(dev) if bias else None, act,
             rs.to(dev) if res else None, N)
    torch.cuda.synchronize()
    ops.MATH_MODE = 0
    err = (C.cpu().double() - ref).abs().max().item() / math.sqrt(K)
    return err

ok = True
for layout in (0, 1, 2):
    for (M, N, K) in [(128, 128, 32), (128, 128, 64), (128, 128, 512), (256, 256, 96), (200, 130, 100), (1280, 512, 512), (33, 511, 256), (2560, 256, 1024), (512, 2048, 2560)]:
        if layout == 2 and (M % 4 or N % 4): continue
        if layout == 1 and N % 4: continue
        if K % 4 and layout == 0: continue
        e1 = run(layout, M, N, K, 1, act=(M % 2 == 0), bias=True, res=(N % 3 == 0))
        e0 = run(layout, M, N, K, 0, act=(M % 2 == 0), bias=True, res=(N % 3 == 0))
        flag = "OK " if e1 < 3e-6 else "BAD"
        ok &= e1 < 3e-6
        print(f"{flag} layout {layout} M{M} N{N} K{K}: tc err/sqrtK {e1:.2e}   ffma {e0:.2e}", flush=True)
print("ALL OK" if ok else "FAILURES")
# FFMA split-K (tiny outputs)
for (l, M, N, K) in [(2, 2, 256, 4352), (2, 6, 256, 6080), (2, 512, 83, 1280), (0, 40, 8, 4096)]:
    e0 = run(l, M, N, K, 0, bias=True)
    print(("OK " if e0 < 3e-6 else "BAD"), "ffma split-K", l, M, N, K, f"{e0:.2e}")
