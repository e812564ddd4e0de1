"""GPU timing of GaussianRasterizer.integrate, product vs the unmodified reference kernels (oracle/_ref), on a
meshing-shaped workload: P Gaussians, PN query points around them. usage: bench_integrate.py P PN W H [reps]"""
import json
import sys

import torch

import conftest  # noqa: F401
import ref_binding
import test_integrate_gpu as T
from ed3dgs_b200.rasterizer import GaussianRasterizer

P, PN, W, H = [int(x) for x in (sys.argv[1:5] if len(sys.argv) > 4 else (300_000, 2_000_000, 802, 550))]
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 5
rs, act, pts = T.make_integrate_case(P, PN, W, H, splat_scale=0.3)
kw = dict(shs=act["shs"], scales=act["scales"], rotations=act["rotations"])


def timed(fn):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        out = fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, out


r = GaussianRasterizer(rs)
ms_mine, mine = timed(lambda: r.integrate(pts, act["means3D"], None, act["opacities"], **kw))
res = dict(P=P, PN=PN, W=W, H=H, ms_product=ms_mine)
if ref_binding.available():
    ref = ref_binding.RefRasterizer()
    ms_ref, (Rn, out) = timed(lambda: ref.integrate(rs, pts, act["means3D"], act["opacities"], **kw))
    res.update(ms_reference=ms_ref, speedup=ms_ref / ms_mine, num_rendered=Rn,
               alpha_max_abs_diff=float((mine[1] - out["alpha_integrated"]).abs().max()),
               alpha_mismatch=int((mine[1] != out["alpha_integrated"]).sum()),
               sdf_max_abs_diff=float((mine[4] - out["point_sdf"]).abs().max()),
               image_max_abs_diff=float((mine[0][:8] - out["color"][:8]).abs().max()),
               max_points_per_pixel=float(out["color"][8].max()),
               points_in_image=int((out["point_sdf"] != -1000).sum()))
print(json.dumps(res))
