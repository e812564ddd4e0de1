"""One warm-up + one measured rasterizer fwd+bwd for ncu launch lists.
Usage: python tests/profile_raster.py [mine|ref] [C2|C3] [splat_scale]"""
import sys

import torch

import conftest  # noqa: F401
import parity_util as pu
from ed3dgs_b200 import synthetic


def main():
    who = sys.argv[1] if len(sys.argv) > 1 else "mine"
    name = sys.argv[2] if len(sys.argv) > 2 else "C3"
    scale = float(sys.argv[3]) if len(sys.argv) > 3 else 0.3
    c = synthetic.CONFIGS[name]
    rs, inp = pu.make_case(P=c["P"], W=c["W"], H=c["H"], sh_degree=c["sh_degree"], splat_scale=scale)
    for it in range(2):
        torch.cuda.synchronize()
        if who == "mine":
            import stage_util as su
            su.product_backward_stages(rs, inp, pu.upstream_grads)
        else:
            pu.run_reference(rs, inp, pu.upstream_grads)
        torch.cuda.synchronize()
    print("done")


if __name__ == "__main__":
    main()
