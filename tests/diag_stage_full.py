"""Stage-wise backward parity at BASELINE's full sizes, several runs each (GPU box):
python tests/diag_stage_full.py [runs] [C3 C4 C5 ...]"""
import sys

import torch
import conftest  # noqa: F401
import parity_util as pu
import stage_util as su
from ed3dgs_b200 import synthetic

runs = int(sys.argv[1]) if len(sys.argv) > 1 else 3
names = sys.argv[2:] or ["C3", "C4", "C5"]
for name in names:
    c = synthetic.CONFIGS[name]
    rs, inp = pu.make_case(P=c["P"], W=c["W"], H=c["H"], sh_degree=c["sh_degree"],
                           focal=c["focal_full"] * c["W"] / c["full_w"])
    for r in range(runs):
        rep = su.stage_stats(rs, inp)
        for st in ("A", "B"):
            for k, v in rep[st].items():
                print(f"{name} run{r} {st} {k:20s} max_rel={v['max_rel']:.3e} l2_rel={v['l2_rel']:.3e} "
                      f"outliers={v['outliers']} l2_trim={v['l2_trim']:.3e} nonfinite_diff={v['nonfinite_diff']} "
                      f"worst={[(i, round(e, 6)) for i, e in v['worst'][:4]]}")
        sys.stdout.flush()
    # end to end, product vs reference and reference vs reference
    for r in range(runs):
        a = pu.run_product(rs, inp, pu.upstream_grads)
        b = pu.run_reference(rs, inp, pu.upstream_grads)
        b2 = pu.run_reference(rs, inp, pu.upstream_grads)
        for k in ("dL_dmeans3D", "dL_dcov3D", "dL_dscales", "dL_drotations"):
            m = su.gaussian_stats(a[3][k], b[3][k])
            n = su.gaussian_stats(b2[3][k], b[3][k])
            print(f"{name} run{r} E2E {k:16s} mine: max_rel={m['max_rel']:.3e} outliers={m['outliers']} "
                  f"l2_trim={m['l2_trim']:.3e} | ref-ref: max_rel={n['max_rel']:.3e} outliers={n['outliers']} "
                  f"l2_trim={n['l2_trim']:.3e}")
        del a, b, b2
        torch.cuda.empty_cache()
        sys.stdout.flush()
