import sys, ctypes
sys.path.insert(0, "/root/repo")
import torch
from grappa_b200 import _lib, ops
_lib.LIB_PATH = "/root/repo/scratch/libtl.so"
# only gemm symbols exist in this debug build: bind manually
h = ctypes.CDLL(_lib.LIB_PATH)
fn = h.grappa_gemm; fn.restype = ctypes.c_int; fn.argtypes = _lib.PROTOTYPES["grappa_gemm"][1]
tl = h.grappa_debug_timeline; tl.restype = ctypes.c_int
dev = "cuda"
for (M, N, K) in [(1536, 256, 256), (6080, 256, 256), (12160, 512, 512), (128, 128, 256), (128, 128, 32)]:
    A = torch.randn(M, K, device=dev); B = torch.randn(N, K, device=dev); C = torch.empty(M, N, device=dev)
    for it in range(3):
        rc = fn(0, M, N, K, A.data_ptr(), K, B.data_ptr(), K, C.data_ptr(), N, None, 0, None, 0, None, 0, 0.0, 0, None, 1, None, 0, None)
        assert rc == 0, h.grappa_last_error()
        torch.cuda.synchronize()
    out = (ctypes.c_ulonglong * 16)()
    tl(out)
    t = [out[i] for i in range(9)]
    names = ["start", "prologue done", "first TMA landed", "first chunk full", "last chunk full", "accum done", "staged+bar", "epilogue done", "all synced"]
    print(f"M{M} N{N} K{K}: " + ", ".join(f"{n}=+{(t[i]-t[0])/1e3:.2f}us" for i, n in enumerate(names)))
