"""Fused activations (csrc/activations.cu) against the reference's own torch statements
(gaussian_renderer/__init__.py:77-83, scene/gaussian_model.py:37-45,594-603), forward and backward.
Floating point: both sides are fp32 elementwise chains, tolerance 2e-6 relative (1 ulp-level differences between
libdevice expf/sqrtf and torch's kernels), written below."""
import pytest
import torch

RTOL, ATOL = 2e-6, 1e-7


def _reference(s, q, o, f):
    """The reference's statements, verbatim semantics."""
    rot = torch.nn.functional.normalize(q)
    if f is None:
        return torch.exp(s), rot, torch.sigmoid(o)
    opacity = torch.sigmoid(o)
    scales = torch.exp(s)
    scales_square = torch.square(scales)
    det1 = scales_square.prod(dim=1)
    scales_after_square = scales_square + torch.square(f)
    det2 = scales_after_square.prod(dim=1)
    coef = torch.sqrt(det1 / det2)
    return torch.sqrt(scales_after_square), rot, opacity * coef[..., None]


def _inputs(P, dev, seed=0):
    g = torch.Generator(device="cpu").manual_seed(seed)
    s = (torch.randn(P, 3, generator=g) * 0.6 - 5.5).to(dev)
    q = torch.randn(P, 4, generator=g).to(dev)
    o = (torch.randn(P, 1, generator=g) * 2.0).to(dev)
    f = (torch.rand(P, 1, generator=g) * 0.01 + 1e-4).to(dev)
    return s, q, o, f


@pytest.mark.gpu
@pytest.mark.parametrize("with_filter", [False, True])
@pytest.mark.parametrize("P", [1, 33, 100003])
def test_activations_match_torch(P, with_filter):
    from ed3dgs_b200.activations import gaussian_activations
    dev = torch.device("cuda:0")
    s, q, o, f = _inputs(P, dev)
    if P > 40:
        q[7] = 0.0                      # below the normalize clamp: y = q / eps, dq = g / eps
        q[9] = torch.tensor([3e-13, 0, 0, 0], device=dev)
    f_in = f if with_filter else None
    a = [t.clone().requires_grad_(True) for t in (s, q, o)]
    b = [t.clone().requires_grad_(True) for t in (s, q, o)]
    out_a = gaussian_activations(a[0], a[1], a[2], f_in)
    out_b = _reference(b[0], b[1], b[2], f_in)
    ups = [torch.randn_like(t) for t in out_b]
    for x, y in zip(out_a, out_b):
        assert x.shape == y.shape
        torch.testing.assert_close(x, y, rtol=RTOL, atol=ATOL)
    torch.autograd.backward(out_a, ups)
    torch.autograd.backward(out_b, ups)
    for x, y in zip(a, b):
        scale = float(y.grad.abs().max()) + 1e-30
        torch.testing.assert_close(x.grad, y.grad, rtol=2e-5, atol=2e-6 * scale)


@pytest.mark.gpu
def test_activations_partial_gradients_and_empty():
    from ed3dgs_b200.activations import gaussian_activations
    dev = torch.device("cuda:0")
    s, q, o, f = _inputs(257, dev, seed=3)
    a = [t.clone().requires_grad_(True) for t in (s, q, o)]
    b = [t.clone().requires_grad_(True) for t in (s, q, o)]
    # only the opacity output carries a gradient: the other upstream gradients arrive as None
    gaussian_activations(a[0], a[1], a[2], f)[2].sum().backward()
    _reference(b[0], b[1], b[2], f)[2].sum().backward()
    # d opacity / d scales is a difference of two nearly equal terms (det1 vs det2 parts): absolute tolerance
    # relative to the largest gradient, as in the test above
    for x, y in ((a[0].grad, b[0].grad), (a[2].grad, b[2].grad)):
        torch.testing.assert_close(x, y, rtol=2e-5, atol=2e-6 * float(y.abs().max()))
    assert float(a[1].grad.abs().max()) == 0.0
    e = [torch.zeros(0, k, device=dev) for k in (3, 4, 1)]
    out = gaussian_activations(e[0], e[1], e[2], None)
    assert [tuple(t.shape) for t in out] == [(0, 3), (0, 4), (0, 1)]


def test_activations_refuse_cpu_tensors():
    from ed3dgs_b200 import _lib
    from ed3dgs_b200.activations import gaussian_activations
    s, q, o, _ = _inputs(4, torch.device("cpu"))
    with pytest.raises(_lib.Ed3dgsError):
        gaussian_activations(s, q, o, None)


@pytest.mark.gpu
@pytest.mark.parametrize("with_filter", [False, True])
def test_render_same_with_and_without_fused_activations(with_filter):
    """The whole step (network -> activations -> rasterizer -> loss -> backward) through renderer.render with
    the fused kernels and with the reference's torch statements: same images, same parameter gradients."""
    from types import SimpleNamespace as NS
    from ed3dgs_b200 import synthetic
    from ed3dgs_b200.renderer import GaussianScene, default_hyper, render, training_loss
    dev = torch.device("cuda:0")
    P, W, H = 6000, 160, 112
    g = synthetic.make_gaussians(P, splat_scale=1.0, device=dev)
    cam = synthetic.make_camera(3, 8, W, H, 8200.0 * W / 3208.0 * 3.0, frame=20, frames=150, device=dev)
    bg = torch.tensor([0.2, 0.1, 0.3], device=dev)
    gt = torch.rand(3, H, W, device=dev, generator=torch.Generator(device=dev).manual_seed(5))
    results = []
    for fused in (True, False):
        torch.manual_seed(1)
        scene = GaussianScene(g, default_hyper()).to(dev)
        scene.fused_activations = fused
        with torch.no_grad():
            scene._deformation.weight.mul_(30.0)
        if with_filter:
            scene.filter_3D = torch.full((P, 1), 2e-3, device=dev)
        pkg = render(cam, scene, NS(debug=False), bg, 0.0, require_coord=False, require_depth=True, cam_no=1,
                     iter=20000, num_down_emb_c=30, num_down_emb_f=30, disable_filter3D=not with_filter)
        training_loss(pkg, gt, view=cam).backward()
        results.append((pkg["render"].detach(), pkg["radii"], scene._scaling.grad.clone(), scene._rotation.grad.clone(),
                        scene._opacity.grad.clone(), scene._embedding.grad.clone()))
    a, b = results
    assert int((a[1] != b[1]).sum()) <= 2           # radii: an ulp in a scale can move a ceil()
    assert float((a[0] - b[0]).abs().max()) < 2e-4
    for x, y in zip(a[2:], b[2:]):
        scale = float(y.abs().max()) + 1e-30
        bad = ((x - y).abs() > 1e-3 * scale + 1e-4 * y.abs()).float().mean()
        assert float(bad) < 2e-3, float(bad)
