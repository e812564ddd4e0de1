"""GPU-vs-GPU baseline of BASELINE.md (B1 + B2), same seeded C3 scene, CUDA events:
  reference side = torch fp32 restatement of deform_network (oracle/deform_oracle.py: the same
  F.linear / F.grid_sample calls as scene/deformation.py -> cuBLAS SGEMM, TF32 off) + the UNMODIFIED
  reference CUDA rasterizer (oracle/_ref), wired through torch autograd like the reference wires it;
  product side  = ed3dgs_b200 render() + backward.
Test infrastructure (uses oracle/): results are committed under profiles/, bench.py does not run it."""
import json
import os
import sys
from types import SimpleNamespace as NS

import torch

import conftest  # noqa: F401
sys.path.insert(0, os.path.join(conftest.ROOT, "oracle"))
sys.path.insert(0, conftest.ROOT)
import deform_oracle  # noqa: E402
import ref_binding  # noqa: E402
import bench  # noqa: E402
from ed3dgs_b200 import synthetic  # noqa: E402
from ed3dgs_b200.renderer import render, training_loss  # noqa: E402


class RefRaster(torch.autograd.Function):
    """The reference's _RasterizeGaussians (diff_gaussian_rasterization/__init__.py:46-174) on the shim."""

    @staticmethod
    def forward(ctx, means3D, sh, opac, scales, rots, rs, tongue, ref):
        R, out = ref.forward(rs, means3D.contiguous(), opac.contiguous(), tongue, shs=sh.contiguous(),
                             scales=scales.contiguous(), rotations=rots.contiguous())
        ctx.ref = ref
        return out["color"], out["depth"], out["mdepth"], out["alpha"], out["normal"]

    @staticmethod
    def backward(ctx, gc, gd, gmd, ga, gn):
        g = ctx.ref.backward(dict(color=gc, depth=gd, mdepth=gmd, alpha=ga, normal=gn))
        return g["dL_dmeans3D"], g["dL_dsh"], g["dL_dopacity"], g["dL_dscales"], g["dL_drotations"], None, None, None


def main():
    torch.backends.cuda.matmul.allow_tf32 = False
    cfgname = sys.argv[1] if len(sys.argv) > 1 else "C3"
    cfg = synthetic.CONFIGS[cfgname]
    dev = torch.device("cuda:0")
    scene, cams = bench.build_scene(cfg, 0.3, dev)
    cam = bench.cam_to(cams[3], dev, False)
    gt = torch.rand(3, cfg["H"], cfg["W"], device=dev)
    bg = torch.zeros(3, device=dev)
    rs = synthetic.raster_settings(cam, cfg["sh_degree"], kernel_size=0.0, require_coord=False, require_depth=True, bg=bg)
    ref = ref_binding.RefRasterizer()
    params = dict(scene.named_parameters())
    sd = {k[len("_deformation."):]: v for k, v in params.items() if k.startswith("_deformation.")}
    P = cfg["P"]

    def ref_step():
        for p in scene.parameters():
            p.grad = None
        time_emb = torch.full((P, 1), cam.time, device=dev)
        shs = torch.cat((scene._features_dc, scene._features_rest), dim=1)
        out = deform_oracle.forward(sd, scene.hyper, 1, 150, scene._xyz, scene._scaling, scene._rotation,
                                    scene._opacity, time_emb, 1, scene._embedding, shs, 20000)
        rots = torch.nn.functional.normalize(out[2])
        scales, opac = torch.exp(out[1]), torch.sigmoid(out[3])
        color, depth, mdepth, alpha, normal = RefRaster.apply(out[0], out[4], opac, scales, rots, rs,
                                                              scene.tongue_class, ref)
        pkg = dict(render=color, expected_depth=depth, median_depth=mdepth, mask=alpha, normal=normal)
        training_loss(pkg, gt, view=cam, fused=False).backward()   # the reference's torch losses

    def our_step():
        for p in scene.parameters():
            p.grad = None
        pkg = render(cam, scene, NS(debug=False), bg, 0.0, require_coord=False, require_depth=True, cam_no=1,
                     iter=20000, num_down_emb_c=30, num_down_emb_f=30)
        training_loss(pkg, gt, view=cam).backward()

    def timeit(fn, warm=3, iters=8):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(iters):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            fn()
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ts.sort()
        return ts[len(ts) // 2]

    res = {"config": cfgname, "P": P}
    res["reference_gpu_ms"] = timeit(ref_step)
    # deformation part of the reference alone (fwd + bwd)
    def ref_mlp():
        for p in scene.parameters():
            p.grad = None
        time_emb = torch.full((P, 1), cam.time, device=dev)
        shs = torch.cat((scene._features_dc, scene._features_rest), dim=1)
        out = deform_oracle.forward(sd, scene.hyper, 1, 150, scene._xyz, scene._scaling, scene._rotation,
                                    scene._opacity, time_emb, 1, scene._embedding, shs, 20000)
        sum((o * 0.37).sum() for o in out[:5]).backward()
    res["reference_mlp_fwd_bwd_ms"] = timeit(ref_mlp)
    torch.cuda.empty_cache()
    torch.cuda.empty_cache()   # the reference arm leaves ~40 GB of differently sized cached blocks behind
    res["product_ms"] = timeit(our_step)
    res["speedup"] = res["reference_gpu_ms"] / res["product_ms"]
    print(json.dumps(res))


if __name__ == "__main__":
    main()
