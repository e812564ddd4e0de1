"""Generates tests/golden/raster_golden.pt on the GPU box from the UNMODIFIED reference CUDA
rasterizer (oracle/_ref/libref_ed3dgs.so, built from /root/reference by oracle/Makefile):
inputs, all 9 outputs, the intermediate integer state and all 8 gradients of small seeded scenes.
Run:  gpurun -- 'cd tests && python golden/make_raster_golden.py'  then copy
gpurun_out/raster_golden.pt to tests/golden/."""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import conftest  # noqa: F401,E402
import parity_util as pu  # noqa: E402
import ref_binding  # noqa: E402

CASES = [
    dict(name="c011", P=700, W=80, H=64, require_coord=False, require_depth=True, kernel_size=0.0),
    dict(name="c000", P=700, W=80, H=64, require_coord=False, require_depth=False, kernel_size=0.0),
    dict(name="c111_ks", P=600, W=70, H=50, require_coord=True, require_depth=True, kernel_size=0.1),
    dict(name="c011_sh0", P=500, W=64, H=64, sh_degree=0, require_coord=False, require_depth=True, kernel_size=0.1),
]


def main():
    recs = []
    for c in CASES:
        c = dict(c)
        name = c.pop("name")
        rs, inp = pu.make_case(splat_scale=1.2, focal=8200.0 * c["W"] / 3208.0 * 3.0, **c)
        R, out, state, grads = pu.run_reference(rs, inp, pu.upstream_grads)
        up = pu.upstream_grads(rs, out)
        cpu = lambda d: {k: (v.detach().cpu() if isinstance(v, torch.Tensor) else v) for k, v in d.items()}  # noqa: E731
        rsd = {k: (v.detach().cpu() if isinstance(v, torch.Tensor) else v) for k, v in rs._asdict().items()}
        keep = ("tiles_touched", "point_offsets", "keys_unsorted", "vals_unsorted", "keys_sorted", "point_list",
                "ranges", "n_contrib", "depths", "means2D")
        recs.append(dict(name=name, case=c, settings=rsd, inputs=cpu({k: v for k, v in inp.items() if v is not None}),
                         R=R, out=cpu(out), state=cpu({k: state[k] for k in keep}), upstream=cpu(up),
                         grads=cpu({k: grads[k] for k in ("dL_dmeans2D", "dL_dcolors", "dL_dopacity", "dL_dmeans3D",
                                                          "dL_dcov3D", "dL_dsh", "dL_dscales", "dL_drotations")})))
        print(name, "R =", R, "visible =", int((out["radii"] > 0).sum()))
    dst = os.path.join(os.environ.get("GRAFT_REPO_ROOT", os.path.dirname(os.path.dirname(HERE))), "gpurun_out")
    os.makedirs(dst, exist_ok=True)
    torch.save(recs, os.path.join(dst, "raster_golden.pt"))
    print("wrote", os.path.join(dst, "raster_golden.pt"), os.path.getsize(os.path.join(dst, "raster_golden.pt")) // 1024, "KiB")


if __name__ == "__main__":
    main()
