"""Decode which shared-memory word the tensor core reads for B(n, k) in MN-major mode."""
import ctypes

import torch

import conftest  # noqa: F401
from ed3dgs_b200 import _lib as L


def main():
    lib = L.lib()
    f = lib.ed3dgs_test_tc_gemm
    f.restype = ctypes.c_int
    vp, ci, cl = ctypes.c_void_p, ctypes.c_int, ctypes.c_long
    f.argtypes = [ci, ci, ci, ci, ci, ci, vp, cl, vp, cl, vp, cl, ci, ci, vp, ci, ci, vp]
    dev = "cuda"
    M, N, K = 128, 128, 16
    n = torch.arange(N, device=dev).view(1, N)
    k = torch.arange(K, device=dev).view(K, 1)
    pos = ((((k // 8) * (N // 4) + n // 4) * 8 + k % 8) * 4 + n % 4).float()   # [K, N] stored layout B[k][n]
    for bmaj, name in ((1, "B MN-major"),):
        for ksel in (0, 1, 5, 8, 13):
            A = torch.zeros(M, K, device=dev)
            A[:, ksel] = 1.0
            C = torch.full((M, N), -1.0, device=dev)
            st = f(0, bmaj, M, N, N, K, A.data_ptr(), K, pos.contiguous().data_ptr(), N, C.data_ptr(), N, 0, -1, None, 0, 2, None)
            L.check(st)
            torch.cuda.synchronize()
            row = C[0].long().tolist()
            exp = pos[ksel].long().tolist()
            print(name, "ksel", ksel, "got[:12]", row[:12], "exp[:12]", exp[:12], "rows_equal", bool((C == C[0:1]).all()))
    # A MN-major probe: D[m][n] with B one-hot (B K-major can't be mixed with A MN in this kernel) -> use MN/MN
    for ksel in (0, 3, 9):
        Apos = pos.clone()                     # A stored [K,128] as A[k][m]
        B = torch.zeros(K, N, device=dev)
        B[ksel, :] = 1.0                       # B(n,k)=1 iff k==ksel  -> D[m][n] = A(m,ksel) for all n (if B works)
        C = torch.zeros(M, N, device=dev)
        st = f(1, 1, M, N, N, K, Apos.contiguous().data_ptr(), 128, B.data_ptr(), N, C.data_ptr(), N, K, -1, None, 4, 2, None)
        L.check(st)
        torch.cuda.synchronize()
        print("A MN-major ksel", ksel, "col0[:12]", C[:12, 0].long().tolist(), "exp", Apos[ksel][:12].long().tolist())


if __name__ == "__main__":
    main()
