"""C2-scale (P=300K) gradient diagnosis: stage-wise parity + worst Gaussian vs reference run-to-run noise."""
import torch
import conftest  # noqa: F401
import parity_util as pu
import stage_util as su

rs, inp = pu.make_case(P=300_000, W=802, H=550)
for k, v in su.stage_report(rs, inp).items():
    print(f"{k:24s} max_rel={v['max_rel']:.3e} l2_rel={v['l2_rel']:.3e}")
a = pu.run_product(rs, inp, pu.upstream_grads)
b = pu.run_reference(rs, inp, pu.upstream_grads)
b2 = pu.run_reference(rs, inp, pu.upstream_grads)
for key in ("dL_dscales", "dL_dcov3D"):
    w = pu.worst_gaussian(a[3], b[3], b2[3], inp, key)
    print(key, w)
    i = w["idx"]
    for k2 in ("dL_dmeans2D", "dL_dopacity", "dL_dmeans3D", "dL_dcov3D"):
        print("   ", k2, "mine", a[3][k2][i].flatten().tolist()[:6], "ref", b[3][k2][i].flatten().tolist()[:6])
    print("    radii", int(a[2]["radii"][i]) if "radii" in a[2] else None, "rot", inp["rotations"][i].tolist())
# distribution: how many Gaussians exceed 1e-3 * scale, product-vs-ref and ref-vs-ref
for key in ("dL_dscales", "dL_dcov3D", "dL_drotations"):
    sc = float(b[3][key].abs().max())
    e1 = (a[3][key] - b[3][key]).abs().flatten(1).max(1).values / sc
    e2 = (b2[3][key] - b[3][key]).abs().flatten(1).max(1).values / sc
    print(key, "count>1e-3: mine", int((e1 > 1e-3).sum()), "ref-ref", int((e2 > 1e-3).sum()),
          "count>1e-4: mine", int((e1 > 1e-4).sum()), "ref-ref", int((e2 > 1e-4).sum()))
