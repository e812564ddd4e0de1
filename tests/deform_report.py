"""Standalone: prints the per-tensor relative errors of the GPU deformation path."""
import sys

import torch

import conftest  # noqa: F401
import test_deform_gpu as T


def main():
    recs = torch.load(T.GOLDEN, weights_only=False)
    recs = list(recs) + [T.oracle_record(N=5000 + 37, kink_free=True), T.oracle_record(N=700, W=64, D=2, it=100, cam_no=None, kink_free=True, no_ds=True, no_dc=True, no_fine_deform=True), T.oracle_record(N=300, W=128, D=1, it=9000, cam_no=1, kink_free=True, no_coarse_deform=True, no_dr=True)]
    for i, rec in enumerate(recs):
        rep = {}
        try:
            T.check(rec, *T.run_product(rec), report=rep)
            status = "PASS"
        except AssertionError as e:
            status = "FAIL " + str(e)[:120]
        worst = sorted(rep.items(), key=lambda kv: -kv[1])[:8]
        print(f"case {i}: {status}")
        for k, v in worst:
            print(f"    {k}: {v:.3e}")
        sys.stdout.flush()


if __name__ == "__main__":
    main()
