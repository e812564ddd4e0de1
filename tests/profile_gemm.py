"""One rows-GEMM launch (M x 128 x 128, optional dbg knobs) for ncu captures."""
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
dbg = int(sys.argv[2]) if len(sys.argv) > 2 else 0
K = n = 128
A = torch.randn(M, K, device="cuda")
B = torch.randn(n, K, device="cuda")
C = torch.zeros(M, n, device="cuda")
bg = torch.zeros(128, device="cuda")
for it in range(2):
    L.check(f(0, 0, M, n, 128, K, A.data_ptr(), K, B.data_ptr(), K, C.data_ptr(), n, 0, -1, bg.data_ptr(), 0, dbg, None))
torch.cuda.synchronize()
print("done")
