"""Standalone GPU parity report (product vs unmodified reference CUDA) -- prints one block per
case. Usage on the GPU box: python tests/parity_report.py [small|full]"""
import json
import sys
import time

import torch

import conftest  # noqa: F401  (sys.path)
import parity_util as pu

CASES_SMALL = [
    dict(name="tiny_011", P=2000, W=128, H=96, require_coord=False, require_depth=True, kernel_size=0.0),
    dict(name="tiny_000", P=2000, W=128, H=96, require_coord=False, require_depth=False, kernel_size=0.0),
    dict(name="tiny_111_ks", P=2000, W=128, H=96, require_coord=True, require_depth=True, kernel_size=0.1),
    dict(name="tiny_101", P=2000, W=100, H=77, require_coord=True, require_depth=False, kernel_size=0.1),
    dict(name="c1_011", P=10000, W=256, H=256, sh_degree=0, require_coord=False, require_depth=True),
    dict(name="mid_011", P=100000, W=802, H=550, require_coord=False, require_depth=True, kernel_size=0.1),
]
CASES_FULL = [
    dict(name="c2_011", P=300000, W=802, H=550, require_coord=False, require_depth=True),
    dict(name="c3_011", P=1000000, W=1604, H=1100, require_coord=False, require_depth=True),
]


def main():
    which = sys.argv[1] if len(sys.argv) > 1 else "small"
    cases = CASES_SMALL + (CASES_FULL if which == "full" else [])
    ok = True
    for c in cases:
        c = dict(c)
        name = c.pop("name")
        rs, inp = pu.make_case(**c)
        t0 = time.time()
        a = pu.run_product(rs, inp, pu.upstream_grads)
        b = pu.run_reference(rs, inp, pu.upstream_grads)
        rep = pu.compare(rs, a, b)
        b2 = pu.run_reference(rs, inp, pu.upstream_grads)
        noise = pu.compare(rs, b2, b)
        rep.update({"REFNOISE_" + k: v for k, v in noise.items() if k.startswith("grad_") and k.endswith("_relerr")})
        import stage_util as su
        for k, v in su.stage_report(rs, inp).items():
            print(f"  STAGE {k}: max_rel={v['max_rel']:.3e} l2_rel={v['l2_rel']:.3e}")
        print(f"=== {name} {c} ({time.time() - t0:.1f}s)")
        print(pu.fmt(rep))
        print("  tiles:", json.dumps(pu.tile_stats(b[2], b[0], c["P"])))
        try:
            pu.assert_report(rep, rs)
            print("  PASS")
        except AssertionError as e:
            ok = False
            print("  FAIL", str(e)[:300])
        sys.stdout.flush()
        del a, b
        torch.cuda.empty_cache()
    print("ALL PASS" if ok else "SOME FAILED")


if __name__ == "__main__":
    main()
