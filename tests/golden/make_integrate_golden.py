"""Generates tests/golden/integrate_golden.pt on the GPU box from the UNMODIFIED reference CUDA kernels
(oracle/_ref/libref_ed3dgs.so: Rasterizer::integrate, rasterizer_impl.cu:580-851): inputs, query points and the
six outputs of GaussianRasterizer.integrate plus invraycov / condition, for small seeded scenes.
Run:  gpurun -- 'cd tests && python golden/make_integrate_golden.py'  then copy gpurun_out/integrate_golden.pt
to tests/golden/."""
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import conftest  # noqa: F401,E402
import ref_binding  # noqa: E402
import test_integrate_gpu as T  # noqa: E402

CASES = [
    dict(name="i_sh3", P=700, PN=3000, W=80, H=64, splat_scale=1.2),
    dict(name="i_ragged", P=500, PN=2500, W=70, H=50, splat_scale=1.5, sh_degree=1),
    dict(name="i_crowded", P=300, PN=6000, W=32, H=32, splat_scale=2.0),  # some pixels hold > 256 points? no: ~6 per px
]


def main():
    recs = []
    for c in CASES:
        c = dict(c)
        name = c.pop("name")
        rs, act, pts = T.make_integrate_case(**c)
        ref = ref_binding.RefRasterizer()
        R, out = ref.integrate(rs, pts, act["means3D"], act["opacities"], shs=act["shs"], scales=act["scales"],
                               rotations=act["rotations"])
        cpu = lambda d: {k: (v.detach().cpu() if isinstance(v, torch.Tensor) else v) for k, v in d.items()}  # noqa: E731
        P = act["means3D"].shape[0]
        out["condition"] = out["condition"][:P]
        recs.append(dict(name=name, case=c, settings=cpu(rs._asdict()), points=pts.cpu(),
                         inputs=cpu({k: act[k] for k in ("means3D", "opacities", "shs", "scales", "rotations")}),
                         R=R, out=cpu(out)))
        inside = out["point_sdf"] != -1000
        print(name, "R =", R, "points in image", int(inside.sum()), "max per pixel", float(out["color"][8].max()),
              "touched", int((out["alpha_integrated"][inside] > 0.01).sum()))
    dst = os.path.join(os.environ.get("GRAFT_REPO_ROOT", os.path.dirname(os.path.dirname(HERE))), "gpurun_out")
    os.makedirs(dst, exist_ok=True)
    torch.save(recs, os.path.join(dst, "integrate_golden.pt"))
    print("wrote", os.path.getsize(os.path.join(dst, "integrate_golden.pt")) // 1024, "KiB")


if __name__ == "__main__":
    main()
