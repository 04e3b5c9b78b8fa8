import sys, ctypes
sys.path.insert(0, "/root/repo")
import torch
from grappa_b200 import _lib
dev = "cuda"
h = ctypes.CDLL("/root/repo/scratch/libtl.so")
fn = h.grappa_gemm; fn.restype = ctypes.c_int; fn.argtypes = _lib.PROTOTYPES["grappa_gemm"][1]
tl = h.grappa_debug_timeline; tl.restype = ctypes.c_int
for (M, N, K) in [(128, 128, 256), (18944, 128, 256), (1536, 256, 256)]:
    A = torch.randn(M, K, device=dev); B = torch.randn(N, K, device=dev); C = torch.empty(M, N, device=dev)
    for it in range(4):
        rc = fn(0, M, N, K, A.data_ptr(), K, B.data_ptr(), K, C.data_ptr(), N, None, 0, None, 0, None, 0, 0.0, 0, None, 1, None, 0, None)
        assert rc == 0
        torch.cuda.synchronize()
    out = (ctypes.c_ulonglong * 16)()
    tl(out)
    t = [out[i] for i in range(13)]
    names = ["start", "prologue", "TMA0", "chunk0", "chunkL", "accum", "staged+bar", "epi done", "synced", "it0", "it1", "it2", "it3"]
    print(f"M{M} N{N} K{K}: " + ", ".join(f"{n}=+{(t[i]-t[0])/1e3:.2f}" for i, n in enumerate(names)))
