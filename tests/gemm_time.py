"""Timing of the rows GEMM (one head-sized problem) with the kernel's timing knobs (dbg bits):
4 = no weight-slab copies, 8 = no activation loads, 16 = no epilogue stores."""
import ctypes, sys
import torch
import conftest  # noqa: F401
from ed3dgs_b200 import _lib as L

lib = L.lib()
f = lib.ed3dgs_test_tc_gemm
f.restype = ctypes.c_int
vp, ci, cl = ctypes.c_void_p, ctypes.c_int, ctypes.c_long
f.argtypes = [ci, ci, ci, ci, ci, ci, vp, cl, vp, cl, vp, cl, ci, ci, vp, ci, ci, vp]
M = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
for (K, n, lda, ldc) in [(128, 128, 128, 128), (128, 128, 640, 640), (32, 128, 32, 128), (128, 32, 128, 32)]:
    A = torch.randn(M, lda, device="cuda")
    B = torch.randn(n, K, device="cuda")
    C = torch.zeros(M, ldc, device="cuda")
    bg = torch.zeros(128, device="cuda")
    for dbg in (0, 16):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        for it in range(3):
            if it == 1:
                ev[0].record()
            L.check(f(0, 0, M, n, 128, K, A.data_ptr(), lda, B.data_ptr(), K, C.data_ptr(), ldc, 0, -1, bg.data_ptr(), 0, dbg, None))
        ev[1].record()
        torch.cuda.synchronize()
        ms = ev[0].elapsed_time(ev[1]) / 2
        print(f"K={K:3d} n={n:3d} lda={lda:3d} ldc={ldc:3d} dbg={dbg:2d}: {ms:7.3f} ms  (incl. prepare)")
