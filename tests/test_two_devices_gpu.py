"""One process driving two GPUs: kernel attributes (dynamic shared memory above 48 KB), SM counts and scratch are
per-device state. The rasterizer and the deformation network run on cuda:0 and then on cuda:1 and must agree."""
import pytest
import torch

from ed3dgs_b200 import synthetic
from ed3dgs_b200.rasterizer import GaussianRasterizer
from ed3dgs_b200.renderer import GaussianScene, default_hyper

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")]


def _raster(dev):
    cam = synthetic.make_camera(1, 8, 320, 200, 700.0, device=dev)
    rs = synthetic.raster_settings(cam, 3, kernel_size=0.1, require_coord=True, require_depth=True)
    g = synthetic.activate(synthetic.make_gaussians(20_000, device=dev))
    means = g["means3D"].clone().requires_grad_(True)
    m2d = torch.zeros_like(means, requires_grad=True)
    out = GaussianRasterizer(rs)(means, m2d, g["opacities"], g["tongue_class"], shs=g["shs"], scales=g["scales"],
                                 rotations=g["rotations"])
    (out[0].sum() + out[4].sum() + out[8].sum()).backward()
    torch.cuda.synchronize(dev)
    return [o.detach().cpu() for o in out], means.grad.cpu()


def _network(dev):
    torch.manual_seed(3)
    N = 5000
    scene = GaussianScene(synthetic.make_gaussians(N, device="cpu"), default_hyper()).to(dev)
    t = torch.full((1, 1), 0.3, device=dev)
    out = scene._deformation(scene._xyz, scene._scaling, scene._rotation, scene._opacity, t, 1, scene, None,
                             scene.get_features, iter=20000, num_down_emb_c=30, num_down_emb_f=30)
    sum(o.sum() for o in out[:5]).backward()
    torch.cuda.synchronize(dev)
    grads = {k: p.grad.cpu() for k, p in scene.named_parameters() if p.grad is not None}
    return [o.detach().cpu() for o in out[:5]], grads


def test_second_device_matches_first():
    a_out, a_g = _raster("cuda:0")
    b_out, b_g = _raster("cuda:1")
    for x, y in zip(a_out, b_out):
        assert torch.equal(x, y)
    assert float((a_g - b_g).abs().max()) <= 1e-4 * float(a_g.abs().max())
    a_out, a_g = _network("cuda:0")
    b_out, b_g = _network("cuda:1")
    for x, y in zip(a_out, b_out):
        assert torch.equal(x, y)
    for k in a_g:
        sc = float(a_g[k].abs().max())
        assert float((a_g[k] - b_g[k]).abs().max()) <= 1e-4 * max(sc, 1e-30), k
    # the KNN front end shares the radix sort
    from ed3dgs_b200 import distCUDA2
    pts = synthetic.make_gaussians(30_000)["xyz"]
    assert torch.equal(distCUDA2(pts.to("cuda:0")).cpu(), distCUDA2(pts.to("cuda:1")).cpu())


def test_overlap_settings_and_static_scratch_are_per_device():
    """The SM reserve of the network's persistent kernels is kept for the device that is current at the call, and the
    built-in scratch of ed3dgs_dist_cuda2_static has one address per device (looked up with that device current)."""
    from ed3dgs_b200 import _lib, distCUDA2
    lib = _lib.lib()
    with torch.cuda.device(0):
        assert lib.ed3dgs_deform_set_sm_reserve(8) == 0
    try:
        with torch.cuda.device(1):
            assert lib.ed3dgs_deform_set_sm_reserve(0) == 0      # cuda:0's setting is not visible here
        with torch.cuda.device(0):
            assert lib.ed3dgs_deform_set_sm_reserve(8) == 8
    finally:
        with torch.cuda.device(0):
            lib.ed3dgs_deform_set_sm_reserve(0)
    pts = synthetic.make_gaussians(30_000)["xyz"]
    for dev in ("cuda:0", "cuda:1", "cuda:0"):
        p = pts.to(dev).contiguous()
        out = torch.empty(p.shape[0], dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            _lib.check(lib.ed3dgs_dist_cuda2_static(p.shape[0], p.data_ptr(), out.data_ptr(), _lib.stream_handle()))
        assert torch.equal(out, distCUDA2(p))
