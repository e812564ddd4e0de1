"""GPU parity of GaussianRasterizer.integrate (GOF point integration for mesh extraction) against the
unmodified reference kernels (oracle/_ref, Rasterizer::integrate rasterizer_impl.cu:580-851).
Bit-exact: radii, point pixel coordinates, the per-pixel point count plane. <= 1e-5 abs: image planes,
integrated alpha / colour; sdf <= 1e-5 relative to the scene depth."""
import pytest
import torch

import ref_binding
from ed3dgs_b200 import synthetic
from ed3dgs_b200.rasterizer import GaussianRasterizer

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(not ref_binding.available(), reason="oracle/_ref not built")

ATOL = 1e-5


def make_integrate_case(P, PN, W, H, splat_scale=1.0, sh_degree=3, seed=synthetic.SEED, cam_k=3, min_scale=3e-4,
                        device="cuda"):
    """Scene + query points shaped like mesh_extract_tetrahedra.py's: most points sit within a few splat radii of
    a Gaussian centre (tetrahedra vertices), some are free in the volume, a few project outside the image.
    Scales are floored at 3e-4 so that every covariance is well-conditioned (min eigenvalue > 1e-8): for an
    ill-conditioned Gaussian the reference stores an uninitialised matrix (forward.cu:200-224)."""
    g = synthetic.make_gaussians(P, seed=seed, splat_scale=splat_scale, device=device)
    act = synthetic.activate(g)
    act["scales"] = act["scales"].clamp_min(min_scale).contiguous()
    cam = synthetic.make_camera(cam_k, 16, W, H, 8200.0 * W / 3208.0, device=device)
    rs = synthetic.raster_settings(cam, sh_degree, bg=torch.tensor([0.1, 0.2, 0.3], device=device))
    gen = torch.Generator(device).manual_seed(seed + 1)
    n_near = int(0.85 * PN)
    owner = torch.randint(0, P, (n_near,), device=device, generator=gen)
    near = act["means3D"][owner] + torch.randn(n_near, 3, device=device, generator=gen) * \
        act["scales"][owner].amax(1, keepdim=True) * 2.0
    free = (torch.rand(PN - n_near, 3, device=device, generator=gen) - 0.5) * 0.7
    pts = torch.cat([near, free], 0).contiguous()
    return rs, act, pts


def run_both(rs, act, pts, **kw):
    r = GaussianRasterizer(rs)
    inputs = dict(shs=act["shs"], scales=act["scales"], rotations=act["rotations"])
    inputs.update(kw)
    mine = r.integrate(pts, act["means3D"], torch.zeros_like(act["means3D"]), act["opacities"], **inputs)
    torch.cuda.synchronize()
    ref = ref_binding.RefRasterizer()
    _, out = ref.integrate(rs, pts, act["means3D"], act["opacities"], **inputs)
    return mine, out


def check(mine, ref, rs):
    color, alpha_i, color_i, coord, sdf, radii = mine
    assert torch.equal(radii, ref["radii"])
    assert torch.equal(coord, ref["point_coordinate"]), "point pixel coordinates must be bit-identical"
    inside = ref["point_sdf"] != -1000.0
    assert 0 < int(inside.sum()) < inside.numel()
    assert torch.equal(color[8], ref["color"][8]), "points-per-pixel plane"
    assert float((color[5]).abs().max()) == 0.0
    for ch in (0, 1, 2, 7):
        d = float((color[ch] - ref["color"][ch]).abs().max())
        assert d <= ATOL, (ch, d)
    for ch in (3, 4, 6):  # depth planes, magnitude ~1
        d = float((color[ch] - ref["color"][ch]).abs().max())
        assert d <= ATOL * max(1.0, float(ref["color"][ch].abs().max())), (ch, d)
    da = (alpha_i - ref["alpha_integrated"]).abs()
    assert float(da.max()) <= ATOL, (float(da.max()), int((da > ATOL).sum()))
    assert float((color_i - ref["color_integrated"]).abs().max()) <= ATOL
    ds = (sdf - ref["point_sdf"]).abs()
    assert float(ds.max()) <= ATOL * max(1.0, float(ref["point_sdf"][inside].abs().max())), float(ds.max())
    # the test is vacuous unless a fair share of the points actually accumulated something
    touched = (ref["alpha_integrated"] > 0.01) & inside
    assert int(touched.sum()) > 0.05 * int(inside.sum()), (int(touched.sum()), int(inside.sum()))
    return dict(alpha=float(da.max()), sdf=float(ds.max()), inside=int(inside.sum()), touched=int(touched.sum()))


@needs_ref
@pytest.mark.parametrize("P,PN,W,H", [(3000, 12000, 160, 120), (20_000, 150_000, 403, 277)])
def test_integrate_matches_reference(P, PN, W, H):
    rs, act, pts = make_integrate_case(P, PN, W, H)
    mine, ref = run_both(rs, act, pts)
    print(check(mine, ref, rs))


@needs_ref
def test_integrate_precomputed_colours_and_crowded_pixels():
    # many points per pixel (> 256 in the busiest ones would take the reference's multi-round path; stay below)
    rs, act, pts = make_integrate_case(2000, 60_000, 64, 48, splat_scale=2.0)
    cols = torch.rand(2000, 3, device="cuda", generator=torch.Generator("cuda").manual_seed(3))
    mine, ref = run_both(rs, act, pts, shs=None, colors_precomp=cols)
    assert float(ref["color"][8].max()) <= 256.0
    print(check(mine, ref, rs))


def test_integrate_empty_inputs_and_defaults():
    rs, act, pts = make_integrate_case(500, 2000, 64, 48)
    r = GaussianRasterizer(rs)
    z3 = torch.zeros(0, 3, device="cuda")
    out = r.integrate(z3, act["means3D"], None, act["opacities"], shs=act["shs"], scales=act["scales"],
                      rotations=act["rotations"])
    assert out[0].shape == (9, 48, 64) and float(out[0].abs().max()) == 0.0 and out[1].numel() == 0
    out = r.integrate(pts, z3, None, torch.zeros(0, 1, device="cuda"), shs=torch.zeros(0, 16, 3, device="cuda"),
                      scales=z3, rotations=torch.zeros(0, 4, device="cuda"))
    assert torch.all(out[1] == 1.0) and torch.all(out[4] == -1000.0) and float(out[2].abs().max()) == 0.0
    # points behind the camera / outside the image keep the reference's fill values
    far = pts.clone()
    far[:, 2] -= 100.0
    out = r.integrate(far, act["means3D"], None, act["opacities"], shs=act["shs"], scales=act["scales"],
                      rotations=act["rotations"])
    assert torch.all(out[1] == 1.0) and torch.all(out[4] == -1000.0) and float(out[3].abs().max()) == 0.0
    assert float(out[0][8].abs().max()) == 0.0 and float(out[0][7].max()) > 0.0


def test_integrate_properties_with_degenerate_gaussians():
    # the default synthetic scene keeps its 1 % needle / pancake Gaussians (ill-conditioned covariances):
    # no reference value exists for those, but the result must stay finite and inside [0, 1)
    g = synthetic.make_gaussians(5000, device="cuda")
    act = synthetic.activate(g)
    cam = synthetic.make_camera(2, 16, 200, 150, 8200.0 * 200 / 3208.0, device="cuda")
    rs = synthetic.raster_settings(cam, 3)
    pts = act["means3D"] + 0.002 * torch.randn_like(act["means3D"])
    out = GaussianRasterizer(rs).integrate(pts, act["means3D"], None, act["opacities"], shs=act["shs"],
                                           scales=act["scales"], rotations=act["rotations"])
    for t in out[:5]:
        assert torch.isfinite(t).all()
    inside = out[4] != -1000.0
    assert float(out[1][inside].min()) >= 0.0 and float(out[1][inside].max()) <= 1.0 + 1e-6
    assert int(out[0][8].sum()) == int(inside.sum())


def test_renderer_integrate_glue_matches_manual_pipeline():
    """ed3dgs_b200.renderer.integrate (mirror of gaussian_renderer.integrate, __init__.py:551-665) = deformation
    network + 3D filter + activations + GaussianRasterizer.integrate."""
    from types import SimpleNamespace as NS
    from ed3dgs_b200.renderer import GaussianScene, default_hyper, integrate
    dev = "cuda"
    g = synthetic.make_gaussians(4000, splat_scale=1.0, device=dev)
    scene = GaussianScene(g, default_hyper()).to(dev)
    scene.filter_3D = torch.full((4000, 1), 3e-4, device=dev)
    cam = synthetic.make_camera(2, 8, 128, 96, 8200.0 * 128 / 3208.0 * 3.0, frame=40, frames=150, device=dev)
    bg = torch.tensor([0.1, 0.2, 0.3], device=dev)
    pts = (g["xyz"] + 0.003 * torch.randn_like(g["xyz"])).contiguous()
    ret = integrate(pts, cam, scene, NS(debug=False), bg, 0.0, loaded_iter=20000, num_down_emb_c=30, num_down_emb_f=30)
    assert set(ret) == {"render", "alpha_integrated", "color_integrated", "point_coordinate", "point_sdf",
                        "visibility_filter", "radii"}
    with torch.no_grad():
        t = torch.full((1, 1), cam.time, device=dev)
        m, s, r, o, sh, _ = scene._deformation(scene._xyz, scene._scaling, scene._rotation, scene._opacity, t, None,
                                               scene, None, scene.get_features, iter=20000, num_down_emb_c=30,
                                               num_down_emb_f=30)
        s, o = scene.apply_scaling_n_opacity_with_3D_filter(opacity=o, scales=s)
        rs = synthetic.raster_settings(cam, 3, bg=bg, require_coord=True)
        out = GaussianRasterizer(rs).integrate(pts, m, None, o, shs=sh, scales=s,
                                               rotations=torch.nn.functional.normalize(r))
    assert torch.equal(ret["alpha_integrated"], out[1]) and torch.equal(ret["render"], out[0])
    inside = ret["point_sdf"] != -1000.0
    assert int(inside.sum()) > 300 and float(ret["alpha_integrated"][inside].max()) > 0.05
