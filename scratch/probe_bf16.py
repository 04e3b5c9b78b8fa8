"""Correctness + speed probe of grappa_gemm_bf16x3 / grappa_split_bf16 (development)."""
import sys
sys.path.insert(0, "/root/repo")
import torch
from grappa_b200 import _lib, ops
L = _lib.lib()
dev = "cuda"
def planes(w):
    N, K = w.shape
    n = int(L.grappa_split_bf16_elems(N, K))
    buf = torch.zeros(n, dtype=torch.int16, device=dev)
    table = torch.tensor([0, N, K, 0], dtype=torch.int64, device=dev)
    mt = ((N + 31) // 32) * ((K + 31) // 32)
    _lib.check(L.grappa_split_bf16(1, table.data_ptr(), mt, w.data_ptr(), buf.data_ptr(), torch.cuda.current_stream().cuda_stream))
    ldk, ldn = (K + 7) // 8 * 8, (N + 7) // 8 * 8
    hi = buf[:N * ldk]; lo = buf[N * ldk:2 * N * ldk]; hiT = buf[2 * N * ldk:2 * N * ldk + K * ldn]; loT = buf[2 * N * ldk + K * ldn:2 * N * ldk + 2 * K * ldn]
    return buf, (hi, lo, ldk), (hiT, loT, ldn)
def run(M, N, K, ws=None, bias=False, act=0, transposed=False):
    torch.manual_seed(M * 7 + N * 3 + K)
    a = torch.randn(M, K, device=dev)
    w = torch.randn(N, K, device=dev) / K ** 0.5
    b = torch.randn(N, device=dev) if bias else None
    if transposed:      # dX-style: C = A . W with W (K, N): planes of W^T come from the hiT / loT block of W's planes
        wt = w.t().contiguous()                       # (K, N) "weight" whose transposed planes are (N, K)
        buf, _, (h, l, ld) = planes(wt)
    else:
        buf, (h, l, ld), _ = planes(w)
    c = torch.empty(M, N, device=dev)
    wsf = ws.numel() if ws is not None else 0
    _lib.check(L.grappa_gemm_bf16x3(M, N, K, a.data_ptr(), K, h.data_ptr(), l.data_ptr(), ld, c.data_ptr(), N,
                                    b.data_ptr() if bias else None, act, None, 0, None, 0, 0.0, 0, None, 1,
                                    ws.data_ptr() if ws is not None else None, wsf, None, 0, torch.cuda.current_stream().cuda_stream))
    ref = a.double() @ w.double().t()
    if bias: ref = ref + b.double()
    if act: ref = torch.nn.functional.elu(ref)
    err = float((c.double() - ref).abs().max() / ref.abs().max())
    return err
ws = torch.empty(8 * 1024 * 1024, device=dev)
for shape in [(128, 128, 64), (128, 128, 128), (256, 128, 512), (1280, 512, 512), (1280, 2048, 512), (1280, 512, 2048), (100, 511, 256), (333, 130, 72), (4352, 256, 1536), (12160, 1536, 512), (200, 256, 4096)]:
    for kw in (dict(), dict(ws=ws, bias=True, act=1), dict(transposed=True)):
        e = run(*shape, **kw)
        print(shape, {k: (v if k != "ws" else "ws") for k, v in kw.items()}, f"rel err {e:.2e}", "OK" if e < 3e-5 else "FAIL", flush=True)
# speed vs the tf32 path
def bench(fn, n=30):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n * 1e3
for (M, N, K) in [(1280, 512, 512), (1280, 2048, 512), (1280, 512, 2048), (12160, 512, 512), (12160, 2048, 512), (12160, 512, 2048), (6080, 256, 2048)]:
    a = torch.randn(M, K, device=dev); w = torch.randn(N, K, device=dev) / K ** 0.5; c = torch.empty(M, N, device=dev)
    buf, (h, l, ld), _ = planes(w); wlo = ops.split_lo(w)
    st = torch.cuda.current_stream().cuda_stream
    t1 = bench(lambda: L.grappa_gemm_bf16x3(M, N, K, a.data_ptr(), K, h.data_ptr(), l.data_ptr(), ld, c.data_ptr(), N, None, 0, None, 0, None, 0, 0.0, 0, None, 1, ws.data_ptr(), ws.numel(), None, 0, st))
    t0 = bench(lambda: L.grappa_gemm_ex(0, M, N, K, a.data_ptr(), K, w.data_ptr(), K, c.data_ptr(), N, None, 0, None, 0, None, 0, 0.0, 0, None, 1, ws.data_ptr(), ws.numel(), None, 0, wlo.data_ptr(), st))
    fl = 2.0 * M * N * K
    print(f"{M}x{N}x{K}: 3xTF32 {t0:.1f} us ({fl / t0 / 1e6:.0f} TF/s)   bf16x3 {t1:.1f} us ({fl / t1 / 1e6:.0f} TF/s)", flush=True)
