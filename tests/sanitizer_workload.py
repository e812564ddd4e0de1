"""Product-only workload for compute-sanitizer ("racecheck / memcheck on the new kernels"): every kernel
of libed3dgs_b200.so once, at sizes the tools finish in minutes. No reference kernels, no oracle: findings are ours.

    compute-sanitizer --tool memcheck|racecheck|synccheck|initcheck python tests/sanitizer_workload.py [part ...]

parts: step (network fused path + activations + rasterizer + losses, forward and backward), layered (network at W = 64,
D = 2), variants (the four rasterizer variants), integrate, knn (distCUDA2 both entry points, K-NN, 3D filter)."""
import sys
from types import SimpleNamespace as NS

import torch
import conftest  # noqa: F401
from ed3dgs_b200 import synthetic

dev = torch.device("cuda:0")


def part_step():
    from ed3dgs_b200.renderer import GaussianScene, default_hyper, render, training_loss
    torch.manual_seed(0)
    P, W, H = 3000, 128, 96
    scene = GaussianScene(synthetic.make_gaussians(P, splat_scale=1.0, device=dev), default_hyper()).to(dev)
    cam = synthetic.make_camera(2, 8, W, H, 8200.0 * W / 3208.0 * 3.0, frame=40, frames=150, device=dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    gt = torch.rand(3, H, W, device=dev)
    for fused_act in (False, True):
        scene.fused_activations = fused_act
        pkg = render(cam, scene, NS(debug=False), bg, 0.0, require_coord=False, require_depth=True, cam_no=1,
                     iter=20000, num_down_emb_c=30, num_down_emb_f=30)
        training_loss(pkg, gt, view=cam).backward()
    torch.cuda.synchronize()


def part_layered():
    import test_deform_gpu as T
    T.run_product(T.oracle_record(kink_free=True, N=3000, W=64, D=2))
    T.run_product(T.oracle_record(kink_free=True, N=2000, W=192, D=1))
    torch.cuda.synchronize()


def part_variants():
    import parity_util as pu
    for coord in (False, True):
        for depth in (False, True):
            rs, inp = pu.make_case(P=2000, W=100, H=77, kernel_size=0.1, require_coord=coord, require_depth=depth)
            pu.run_product(rs, inp, pu.upstream_grads)


def part_integrate():
    from test_integrate_gpu import make_integrate_case
    from ed3dgs_b200.rasterizer import GaussianRasterizer
    rs, act, pts = make_integrate_case(2000, 5000, 128, 96)
    GaussianRasterizer(rs).integrate(pts, act["means3D"], torch.zeros_like(act["means3D"]), act["opacities"],
                                     shs=act["shs"], scales=act["scales"], rotations=act["rotations"])
    GaussianRasterizer(rs).markVisible(act["means3D"])
    torch.cuda.synchronize()


def part_knn():
    from ed3dgs_b200 import _lib, distCUDA2
    from ed3dgs_b200.extra_utils import knn_gpu
    from ed3dgs_b200.filter3d import compute_3D_filter
    pts = synthetic.make_gaussians(6000, device=dev)["xyz"].contiguous()
    a = distCUDA2(pts)
    b = torch.empty_like(a)
    _lib.check(_lib.lib().ed3dgs_dist_cuda2_static(pts.shape[0], pts.data_ptr(), b.data_ptr(), _lib.stream_handle()))
    knn_gpu(pts, 20)
    cams = [synthetic.make_camera(k, 8, 128, 96, 8200.0 * 128 / 3208.0, device=dev) for k in range(8)]
    compute_3D_filter(pts, cams)
    torch.cuda.synchronize()
    assert torch.equal(a, b)


PARTS = dict(step=part_step, layered=part_layered, variants=part_variants, integrate=part_integrate, knn=part_knn)
for name in (sys.argv[1:] or list(PARTS)):
    PARTS[name]()
    print("done", name, flush=True)
print("SANITIZER_WORKLOAD_OK")
