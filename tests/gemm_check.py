"""GPU check of the 3xTF32 tcgen05 GEMM building block against fp64 torch.matmul."""
import ctypes
import sys

import torch

import conftest  # noqa: F401
from ed3dgs_b200 import _lib as L


def run(amaj, bmaj, M, n_valid, NT, K, dbg=0, k_per_cta=0, ones_col=-1, epilogue=0):
    lib = L.lib()
    f = lib.ed3dgs_test_tc_gemm
    f.restype = ctypes.c_int
    vp, ci, cl = ctypes.c_void_p, ctypes.c_int, ctypes.c_long
    f.argtypes = [ci, ci, ci, ci, ci, ci, vp, cl, vp, cl, vp, cl, ci, ci, vp, ci, ci, vp]
    g = torch.Generator("cuda").manual_seed(1)
    dev = "cuda"
    if amaj == 0:   # A[M, K] row-major
        A = torch.randn(M, K, device=dev, generator=g)
        Aop = A
        lda = K
    else:           # A stored [K, 128]; D rows = features (M <= 128)
        A = torch.randn(K, 128, device=dev, generator=g)
        Aop = A.t()[:M]
        lda = 128
    if bmaj == 0:   # B[n_valid, K]
        B = torch.randn(n_valid, K, device=dev, generator=g)
        Bop = B
        ldb = K
    else:           # B stored [K, ldb>=n_valid]
        ldb = ((n_valid + 3) // 4) * 4
        B = torch.randn(K, ldb, device=dev, generator=g)
        Bop = B.t()[:n_valid]
    ref = (Aop.double() @ Bop.double().t())
    bias_grad = torch.zeros(128, device=dev)
    if amaj == 0:
        C = torch.full((M, n_valid), 7.0, device=dev)
        ldc = n_valid
    else:
        C = torch.zeros(M, n_valid, device=dev) if epilogue == 4 else torch.zeros(n_valid, M, device=dev)
        ldc = C.shape[1]
    st = f(amaj, bmaj, M, n_valid, NT, K, A.data_ptr(), lda, B.data_ptr(), ldb, C.data_ptr(), ldc, k_per_cta,
           ones_col, bias_grad.data_ptr(), epilogue, dbg, None)
    L.check(st)
    torch.cuda.synchronize()
    out = C if (amaj == 0 or epilogue == 4) else C.t()
    err = float((out.double() - ref).abs().max() / ref.abs().max())
    extra = ""
    if ones_col >= 0:
        bsum = Aop.double().sum(1)
        extra = f" ones_err={float((bias_grad[:M].double() - bsum).abs().max() / bsum.abs().max()):.2e}"
    return err, extra


def main():
    cases = [
        ("KK 128x128x32", dict(amaj=0, bmaj=0, M=128, n_valid=128, NT=128, K=32)),
        ("KK 300x128x128", dict(amaj=0, bmaj=0, M=300, n_valid=128, NT=128, K=128)),
        ("KK 256x3(16)x128", dict(amaj=0, bmaj=0, M=256, n_valid=3, NT=16, K=128)),
        ("KK 256x48x128", dict(amaj=0, bmaj=0, M=256, n_valid=48, NT=48, K=128)),
        ("KMN 256x128x8", dict(amaj=0, bmaj=1, M=256, n_valid=128, NT=128, K=8)),
        ("KMN 256x128x48", dict(amaj=0, bmaj=1, M=256, n_valid=128, NT=128, K=48)),
        ("KMN 1000x32x128", dict(amaj=0, bmaj=1, M=1000, n_valid=32, NT=32, K=128)),
        ("KMN 512x128x640", dict(amaj=0, bmaj=1, M=512, n_valid=128, NT=128, K=640)),
        ("MNMN 128x128 K=5000", dict(amaj=1, bmaj=1, M=128, n_valid=128, NT=144, K=5000, k_per_cta=2048, epilogue=4)),
        ("MNMN 128x32 K=3000 T", dict(amaj=1, bmaj=1, M=128, n_valid=32, NT=48, K=3000, k_per_cta=1024, epilogue=5)),
        ("MNMN 128x8 K=777", dict(amaj=1, bmaj=1, M=128, n_valid=8, NT=16, K=777, k_per_cta=512, epilogue=5)),
    ]
    for dbg in (0, 1, 2):
        print(f"--- dbg={dbg}")
        for name, kw in cases:
            try:
                err, extra = run(dbg=dbg, **kw)
                print(f"{name:28s} rel_err={err:.3e}{extra}")
            except Exception as e:  # noqa: BLE001
                print(f"{name:28s} EXC {e}")
                return
        sys.stdout.flush()


if __name__ == "__main__":
    main()
