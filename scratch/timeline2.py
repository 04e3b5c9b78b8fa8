import sys, ctypes
sys.path.insert(0, "/root/repo")
import torch
from grappa_b200 import _lib
dev = "cuda"
for libname in sys.argv[1:]:
    h = ctypes.CDLL(libname)
    fn = h.grappa_gemm; fn.restype = ctypes.c_int; fn.argtypes = _lib.PROTOTYPES["grappa_gemm"][1]
    tl = h.grappa_debug_timeline; tl.restype = ctypes.c_int
    for (M, N, K) in [(128, 128, 2048), (18944, 128, 2048)]:
        A = torch.randn(M, K, device=dev); B = torch.randn(N, K, device=dev); C = torch.empty(M, N, device=dev)
        for it in range(3):
            rc = fn(0, M, N, K, A.data_ptr(), K, B.data_ptr(), K, C.data_ptr(), N, None, 0, None, 0, None, 0, 0.0, 0, None, 1, None, 0, None)
            assert rc == 0
            torch.cuda.synchronize()
        out = (ctypes.c_ulonglong * 16)()
        tl(out)
        t = [out[i] for i in range(9)]
        nkb = K // 32
        print(f"{libname.split('/')[-1]:>24} M{M} K{K}: first chunk +{(t[3]-t[0])/1e3:.2f}us last chunk +{(t[4]-t[0])/1e3:.2f}us -> {(t[4]-t[3])/1e3/(nkb-4):.3f} us/k-block; done +{(t[8]-t[0])/1e3:.2f}")
