"""Image losses (utils/loss_utils.py:24-72, utils/graphics_utils.py:240-267, train.py:195-203,249-267): the CPU
oracle against goldens minted from the reference's own functions, the fused CUDA kernels against both.
Stated tolerance: loss values <= 1e-5 abs, gradients <= 1e-4 relative to the gradient's max magnitude."""
import os
import sys
from types import SimpleNamespace as NS

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import loss_oracle  # noqa: E402
from ed3dgs_b200 import synthetic  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden", "loss_golden.pt")
VAL_ATOL, GRAD_RTOL = 1e-5, 1e-4


def _view(rec):
    return synthetic.make_camera(1, 4, rec["W"], rec["H"], 8200.0 * rec["W"] / 3208.0)


def _rel(a, b):
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("idx", [0, 1, 2])
def test_oracle_matches_reference_golden(idx):
    rec = torch.load(GOLDEN, weights_only=False)[idx]
    img = rec["img"].clone().requires_grad_(True)
    loss = loss_oracle.l1_loss(img, rec["gt"]) + rec["lam"] * (1 - loss_oracle.ssim(img, rec["gt"])) / 2
    (d,) = torch.autograd.grad(loss, img)
    assert abs(float(loss) - float(rec["loss"])) <= 1e-7
    assert _rel(d, rec["d_img"]) <= 1e-6
    de, dm, nr = (rec[k].clone().requires_grad_(True) for k in ("depth_e", "depth_m", "normal"))
    dn = loss_oracle.depth_double_to_normal(_view(rec), de, dm)
    assert float((dn - rec["normals_from_depth"]).abs().max()) <= 1e-6
    dnl = loss_oracle.depth_normal_loss(_view(rec), de, dm, nr)
    g = torch.autograd.grad(dnl, (de, dm, nr))
    assert abs(float(dnl) - float(rec["depth_normal_loss"])) <= 1e-7
    for a, k in zip(g, ("d_depth_e", "d_depth_m", "d_normal")):
        assert _rel(a, rec[k]) <= 1e-5, k


@pytest.mark.gpu
@pytest.mark.parametrize("idx", [0, 1, 2])
def test_gpu_losses_match_golden(idx):
    from ed3dgs_b200 import loss_utils as L
    rec = torch.load(GOLDEN, weights_only=False)[idx]
    img = rec["img"].cuda().requires_grad_(True)
    gt = rec["gt"].cuda()
    loss, l1, s = L.photometric_loss(img, gt, rec["lam"])
    loss.backward()
    assert abs(float(l1) - float(rec["Ll1"])) <= VAL_ATOL and abs(float(s) - float(rec["ssim"])) <= VAL_ATOL
    assert abs(float(loss) - float(rec["loss"])) <= VAL_ATOL
    assert _rel(img.grad.cpu(), rec["d_img"]) <= GRAD_RTOL
    # the reference's separate entry points
    img2 = rec["img"].cuda().requires_grad_(True)
    s2, smap = L.ssim(img2, gt)
    total = L.l1_loss(img2, gt) + rec["lam"] * (1 - s2) / 2
    total.backward()
    assert smap is None and _rel(img2.grad.cpu(), rec["d_img"]) <= GRAD_RTOL
    view = _view(rec)
    de, dm, nr = (rec[k].cuda().requires_grad_(True) for k in ("depth_e", "depth_m", "normal"))
    dnl = L.depth_normal_loss(view, de, dm, nr)
    dnl.backward()
    assert abs(float(dnl) - float(rec["depth_normal_loss"])) <= VAL_ATOL
    assert _rel(de.grad.cpu(), rec["d_depth_e"]) <= GRAD_RTOL
    assert _rel(dm.grad.cpu(), rec["d_depth_m"]) <= GRAD_RTOL
    assert _rel(nr.grad.cpu(), rec["d_normal"]) <= GRAD_RTOL
    dn = L.depth_double_to_normal(view, rec["depth_e"].cuda(), rec["depth_m"].cuda())
    assert float((dn.cpu() - rec["normals_from_depth"]).abs().max()) <= 1e-5
    # differentiable like the reference's (train.py:253-266 back-propagates through it): a seeded upstream gradient
    # on the [2,3,H,W] normals against autograd through the oracle's torch statements
    up = torch.randn(dn.shape, generator=torch.Generator().manual_seed(idx + 1))
    a = [rec[k].clone().cuda().requires_grad_(True) for k in ("depth_e", "depth_m")]
    b = [rec[k].clone().requires_grad_(True) for k in ("depth_e", "depth_m")]
    (L.depth_double_to_normal(view, *a) * up.cuda()).sum().backward()
    (loss_oracle.depth_double_to_normal(view, *b) * up).sum().backward()
    for x, y in zip(a, b):
        assert x.grad is not None and _rel(x.grad.cpu(), y.grad) <= GRAD_RTOL
    # the regulariser built from it, as train.py writes it, equals the fused depth_normal_loss
    c = [rec[k].clone().cuda().requires_grad_(True) for k in ("depth_e", "depth_m", "normal")]
    nrm = L.depth_double_to_normal(view, c[0], c[1])
    err = 1 - (c[2].unsqueeze(0) * nrm).sum(dim=1)
    (0.4 * err[0].mean() + 0.6 * err[1].mean()).backward()
    for x, k in zip(c, ("d_depth_e", "d_depth_m", "d_normal")):
        assert _rel(x.grad.cpu(), rec[k]) <= GRAD_RTOL, k


@pytest.mark.gpu
def test_gpu_losses_match_oracle_at_full_resolution():
    from ed3dgs_b200 import loss_utils as L
    H, W = 1100, 1604
    g = torch.Generator().manual_seed(9)
    gt = torch.rand(3, H, W, generator=g)
    img0 = (gt + 0.1 * torch.randn(3, H, W, generator=g)).clamp(0, 1)
    img = img0.cuda().requires_grad_(True)
    loss, _, _ = L.photometric_loss(img, gt.cuda(), 0.2)
    (2.5 * loss).backward()
    ref_img = img0.cuda().requires_grad_(True)     # the oracle is device-agnostic torch: run it on the GPU for speed
    ref = loss_oracle.l1_loss(ref_img, gt.cuda()) + 0.2 * (1 - loss_oracle.ssim(ref_img, gt.cuda())) / 2
    (2.5 * ref).backward()
    assert abs(float(loss) - float(ref)) <= VAL_ATOL
    assert _rel(img.grad, ref_img.grad) <= GRAD_RTOL
    view = synthetic.make_camera(2, 16, W, H, 8200.0 * W / 3208.0)
    yy, xx = torch.meshgrid(torch.arange(H).float(), torch.arange(W).float(), indexing="ij")
    base = 1.0 + 1e-4 * xx + 2e-4 * yy
    d0 = (base + 1e-3 * torch.rand(H, W, generator=g))[None].cuda()
    d1 = (base + 1e-3 * torch.rand(H, W, generator=g))[None].cuda()
    n0 = torch.nn.functional.normalize(torch.randn(3, H, W, generator=g), dim=0).cuda()
    a = [t.clone().requires_grad_(True) for t in (d0, d1, n0)]
    b = [t.clone().requires_grad_(True) for t in (d0, d1, n0)]
    la = L.depth_normal_loss(view, *a)
    la.backward()
    lb = loss_oracle.depth_normal_loss(view, *b)
    lb.backward()
    assert abs(float(la) - float(lb)) <= VAL_ATOL
    for x, y in zip(a, b):
        assert _rel(x.grad, y.grad) <= GRAD_RTOL


def test_losses_refuse_cpu_tensors():
    from ed3dgs_b200 import Ed3dgsError
    from ed3dgs_b200 import loss_utils as L
    with pytest.raises(Ed3dgsError):
        L.ssim(torch.rand(3, 8, 8), torch.rand(3, 8, 8))
    with pytest.raises(Ed3dgsError):
        L.depth_normal_loss(NS(image_width=8, image_height=8, FoVx=0.5, FoVy=0.5), torch.rand(1, 8, 8),
                            torch.rand(1, 8, 8), torch.rand(3, 8, 8))


@pytest.mark.gpu
@pytest.mark.parametrize("reg,with_view", [(True, True), (True, False), (False, True)])
def test_step_loss_node_equals_the_torch_composition(reg, with_view):
    """renderer.training_loss(fused=True) is ONE autograd node (loss_utils._StepLoss); its value and the gradients it
    sends to every rasterizer output must equal the reference's torch composition of the same terms
    (training_loss(fused=False): utils/loss_utils.py:24-72, utils/graphics_utils.py:240-267, train.py:195-271)."""
    from ed3dgs_b200.renderer import training_loss
    H, W = 130, 171
    g = torch.Generator(device="cuda").manual_seed(4)
    cam = NS(image_width=W, image_height=H, FoVx=0.6, FoVy=0.47)
    gt = torch.rand(3, H, W, device="cuda", generator=g)

    def make():
        gg = torch.Generator(device="cuda").manual_seed(9)
        pkg = {"render": torch.rand(3, H, W, device="cuda", generator=gg),
               "expected_depth": 1.0 + torch.rand(1, H, W, device="cuda", generator=gg),
               "median_depth": 1.0 + torch.rand(1, H, W, device="cuda", generator=gg),
               "normal": torch.nn.functional.normalize(torch.randn(3, H, W, device="cuda", generator=gg), dim=0),
               "mask": torch.rand(1, H, W, device="cuda", generator=gg)}
        return {k: v.requires_grad_(True) for k, v in pkg.items()}

    a, b = make(), make()
    view = cam if with_view else None
    la = training_loss(a, gt, view=view, reg=reg, fused=True)
    lb = training_loss(b, gt, view=view, reg=reg, fused=False)
    assert abs(float(la) - float(lb)) <= 1e-5 * max(1.0, abs(float(lb)))
    (la * 1.7).backward()
    (lb * 1.7).backward()
    for k in a:
        if b[k].grad is None:
            assert a[k].grad is None or float(a[k].grad.abs().max()) == 0.0, k
            continue
        scale = float(b[k].grad.abs().max())
        assert float((a[k].grad - b[k].grad).abs().max()) <= 1e-4 * scale + 1e-12, k
