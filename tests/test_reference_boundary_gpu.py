"""The boundary exactly as the reference's render() drives it (gaussian_renderer/__init__.py:28-45,84): settings
scalars as 0-dim CUDA tensors, time as a [N, 1] repeat, the torch activation statements between the network and the
rasterizer -- compared with ed3dgs_b200.renderer.render (python scalars, (1, 1) time, fused activations), plus the
hidden-sync audit SURVEY section 7 asks for."""
import math
from types import SimpleNamespace as NS

import pytest
import torch

import parity_util as pu
import ref_binding
from ed3dgs_b200 import synthetic
from ed3dgs_b200.rasterizer import GaussianRasterizationSettings, GaussianRasterizer, host_scalars
from ed3dgs_b200.renderer import GaussianScene, default_hyper, render, training_loss

pytestmark = pytest.mark.gpu
PIPE = NS(debug=False, compute_cov3D_python=False, convert_SHs_python=False)


def _scene(P, seed=5):
    """Near-isotropic splats: this file is about argument forms, and the needle Gaussians of the default synthetic
    scene make the scale / rotation gradients (and every network weight they feed) move by more than 1e-4 between
    any two runs of the same code (atomics in arrival order through an ill-conditioned inverse, parity_util.py)."""
    torch.manual_seed(seed)
    g = synthetic.make_gaussians(P, device="cuda")
    g["scaling"] = g["scaling"].mean(1, keepdim=True) + 0.1 * torch.randn(P, 3, device="cuda")
    return GaussianScene(g, default_hyper()).cuda()


def reference_style_render(viewpoint_camera, pc, pipe, bg_color, kernel_size, scaling_modifier=1.0, require_coord=True,
                           require_depth=True, cam_no=None, iter=None, num_down_emb_c=5, num_down_emb_f=5):
    """The statements of the reference's render() that touch the drop-in modules, in its order and with its argument
    forms (gaussian_renderer/__init__.py:17-107); written out here because /root/reference is absent on the GPU box."""
    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True, device="cuda") + 0
    screenspace_points.retain_grad()
    means3D = pc.get_xyz
    tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
    raster_settings = GaussianRasterizationSettings(
        image_height=torch.tensor(viewpoint_camera.image_height).cuda(),
        image_width=torch.tensor(viewpoint_camera.image_width).cuda(),
        tanfovx=torch.tensor(tanfovx).cuda(),
        tanfovy=torch.tensor(tanfovy).cuda(),
        bg=bg_color.cuda(),
        scale_modifier=torch.tensor(scaling_modifier).cuda(),
        viewmatrix=viewpoint_camera.world_view_transform.cuda(),
        projmatrix=viewpoint_camera.full_proj_transform.cuda(),
        sh_degree=torch.tensor(pc.active_sh_degree).cuda(),
        campos=viewpoint_camera.camera_center.cuda(),
        prefiltered=False,
        debug=pipe.debug,
        kernel_size=kernel_size,
        require_coord=require_coord,
        require_depth=require_depth,
    )
    time = torch.tensor(viewpoint_camera.time).to(means3D.device).repeat(means3D.shape[0], 1)
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)
    means2D = screenspace_points
    opacity = pc._opacity
    shs = pc.get_features
    scales = pc._scaling
    rotations = pc._rotation
    means3D_final, scales_final, rotations_final, opacity_final, shs_final, extras = pc._deformation(
        means3D, scales, rotations, opacity, time, cam_no, pc, None, shs, iter=iter, num_down_emb_c=num_down_emb_c,
        num_down_emb_f=num_down_emb_f)
    rotations_final = pc.rotation_activation(rotations_final)
    scales_final = pc.scaling_activation(scales_final)
    opacity = pc.opacity_activation(opacity_final)
    outputs = rasterizer(means3D=means3D_final, means2D=means2D, shs=shs_final, colors_precomp=None,
                         opacities=opacity, tongue_class=pc.tongue_class, scales=scales_final,
                         rotations=rotations_final, cov3D_precomp=None)
    assert len(outputs) == 9
    names = ["render", "radii", "expected_coord", "median_coord", "expected_depth", "median_depth", "mask",
             "tongue_mask", "normal"]
    pkg = dict(zip(names, outputs))
    pkg["viewspace_points"] = means2D
    pkg["deformed_gaussian_positions"] = means3D_final
    pkg["sh_coefs_final"] = shs_final
    return pkg


def _weighted(pkg, H, W):
    gen = torch.Generator("cuda").manual_seed(11)
    loss = 0
    for k, c in (("render", 3), ("expected_depth", 1), ("median_depth", 1), ("mask", 1), ("normal", 3),
                 ("expected_coord", 3), ("median_coord", 3)):
        loss = loss + (pkg[k] * torch.rand(c, H, W, device="cuda", generator=gen)).sum() / (H * W)
    return loss


@pytest.mark.parametrize("coord", [False, True])
def test_reference_style_call_equals_renderer(coord):
    P, cam = 60_000, synthetic.make_camera(3, 16, 402, 275, 8200.0 * 402 / 3208.0, frame=37, frames=150, device="cuda")
    pc = _scene(P)
    bg = torch.tensor([0.1, 0.2, 0.3], device="cuda")
    kw = dict(require_coord=coord, require_depth=True, cam_no=3, iter=12_000, num_down_emb_c=30, num_down_emb_f=30)
    grads = []
    pkgs = []
    for mode in ("reference", "mirror_torch_activations", "mirror"):
        pc.zero_grad(set_to_none=True)
        pc.fused_activations = mode == "mirror"
        fn = reference_style_render if mode == "reference" else render
        pkg = fn(cam, pc, PIPE, bg, 0.1 if coord else 0.0, **kw)
        _weighted(pkg, 275, 402).backward()
        torch.cuda.synchronize()
        pkgs.append(pkg)
        grads.append({k: p.grad.clone() for k, p in pc.named_parameters() if p.grad is not None})
        assert pkg["viewspace_points"].grad is not None and pkg["viewspace_points"].grad.shape == (P, 3)
    ref, mir_t, mir = pkgs
    assert int((ref["radii"] > 0).sum()) > P // 4
    for k in ("render", "expected_depth", "median_depth", "mask", "normal", "expected_coord", "median_coord", "radii",
              "tongue_mask"):
        # same kernels on the same values: the 0-dim tensors and the [N, 1] time must change nothing at all
        assert torch.equal(ref[k], mir_t[k]), k
        # fused activations (csrc/activations.cu): exp / sigmoid / normalize differ from torch's in the last ulp, and
        # the needle Gaussians of this scene amplify that (ill-conditioned 2D covariance inverse): the images agree
        # to 1e-3 everywhere and to 1e-5 on all but a few percent of the pixels
        if k == "radii":
            assert int((ref[k] != mir[k]).sum()) <= 4
        else:
            d = (ref[k] - mir[k]).abs()
            assert float(d.max()) <= 1e-3 and float((d > 1e-5).float().mean()) <= 0.05, (k, float(d.max()))
    assert set(grads[0]) == set(grads[1]) == set(grads[2])
    def rows_off(a, b, tol):
        """Fraction of rows (Gaussians / weight rows) whose max error exceeds tol * the tensor's max magnitude."""
        e = (a - b).abs()
        e = e.flatten(1).amax(1) if e.dim() >= 2 else e
        return float((e > tol * float(b.abs().max())).float().mean())

    for k in grads[0]:
        assert float(grads[0][k].abs().max()) > 0, k
        # same values in, atomics in arrival order out
        if grads[0][k].shape[0] == P:
            assert rows_off(grads[1][k], grads[0][k], 1e-4) <= 1e-3, k
            assert rows_off(grads[2][k], grads[0][k], 2e-4) <= 2e-3, k
        else:
            sc = float(grads[0][k].abs().max())
            assert float((grads[1][k] - grads[0][k]).abs().max()) <= 1e-4 * sc, k
            assert float((grads[2][k] - grads[0][k]).abs().max()) <= 3e-4 * sc, k


def test_compute_cov3d_python_flag_behaves_like_the_reference():
    """With pipe.compute_cov3D_python the reference hands scales = rotations = None to the deformation network,
    which indexes them (deformation.py:107): TypeError there, TypeError here."""
    pc = _scene(500)
    cam = synthetic.make_camera(0, 4, 64, 48, 200.0, device="cuda")
    pipe = NS(debug=False, compute_cov3D_python=True, convert_SHs_python=False)
    with pytest.raises(TypeError):
        render(cam, pc, pipe, torch.zeros(3, device="cuda"), 0.0, iter=100)
    cov = pc.get_covariance(1.0)
    assert cov.shape == (500, 6) and bool((cov[:, [0, 3, 5]] > 0).all())


@pytest.mark.skipif(not ref_binding.available(), reason="oracle/_ref not built")
def test_cov3d_precomp_with_device_scalar_settings_vs_reference():
    """cov3D_precomp (what get_covariance feeds, gaussian_renderer/__init__.py:66-67) through the autograd node with
    0-dim CUDA settings, against the reference kernels on the same covariance."""
    rs, inp = pu.make_case(P=40_000, W=402, H=275, precomp_cov=True)
    dev_rs = rs._replace(image_height=torch.tensor(275).cuda(), image_width=torch.tensor(402).cuda(),
                         tanfovx=torch.tensor(rs.tanfovx).cuda(), tanfovy=torch.tensor(rs.tanfovy).cuda(),
                         scale_modifier=torch.tensor(1.0).cuda(), sh_degree=torch.tensor(3).cuda())
    hs = host_scalars(dev_rs)
    # torch.tensor(python float) is float32, exactly what the reference's .item<float>() sees
    assert hs.image_height == 275 and hs.sh_degree == 3 and hs.tanfovx == float(torch.tensor(rs.tanfovx))
    leaves = {k: inp[k].clone().requires_grad_(True) for k in ("means3D", "opacities", "shs", "cov3D_precomp")}
    m2d = torch.zeros_like(leaves["means3D"], requires_grad=True)
    out = GaussianRasterizer(dev_rs)(leaves["means3D"], m2d, leaves["opacities"], inp["tongue_class"],
                                     shs=leaves["shs"], cov3D_precomp=leaves["cov3D_precomp"])
    color, radii, coord, mcoord, depth, mdepth, alpha, tongue, normal = out
    rs32 = rs._replace(tanfovx=hs.tanfovx, tanfovy=hs.tanfovy)
    Rb, ob, sb, gb = pu.run_reference(rs32, inp, pu.upstream_grads)
    assert torch.equal(radii, ob["radii"])
    for k, v in (("color", color), ("depth", depth), ("mdepth", mdepth), ("alpha", alpha), ("normal", normal)):
        assert float((v - ob[k]).abs().max()) <= pu.RENDER_ATOL, k
    g = pu.upstream_grads(rs32, dict(color=color))
    torch.autograd.backward([color, depth, mdepth, alpha, normal],
                            [g["color"], g["depth"], g["mdepth"], g["alpha"], g["normal"]])
    for mine, ref in ((leaves["means3D"].grad, gb["dL_dmeans3D"]), (leaves["opacities"].grad, gb["dL_dopacity"]),
                      (leaves["shs"].grad, gb["dL_dsh"]), (m2d.grad, gb["dL_dmeans2D"])):
        sc = float(ref.abs().max())
        assert float((mine - ref).abs().max()) <= pu.GRAD_RTOL * sc
    mine, ref = leaves["cov3D_precomp"].grad, gb["dL_dcov3D"]
    e = (mine - ref).abs().flatten(1).max(1).values / float(ref.abs().max())
    assert int((e > pu.GRAD_RTOL).sum()) <= 8


def test_no_hidden_synchronisation_in_a_training_step():
    """One training step through the public API with torch's sync detector armed. The only blocking point of the
    step is the rasterizer's read-back of the instance count R inside the library (the reference's
    rasterizer_impl.cu:359), which torch does not see; anything torch-side (.item(), .cpu(), nonzero, a python
    bool of a CUDA tensor) raises here."""
    P = 30_000
    pc = _scene(P)
    cam = synthetic.make_camera(2, 16, 402, 275, 8200.0 * 402 / 3208.0, frame=3, frames=150, device="cuda")
    bg = torch.zeros(3, device="cuda")
    gt = torch.rand(3, 275, 402, device="cuda")

    def step(fn):
        pc.zero_grad(set_to_none=True)
        pkg = fn(cam, pc, PIPE, bg, 0.0, require_coord=False, require_depth=True, cam_no=2, iter=20_000,
                 num_down_emb_c=30, num_down_emb_f=30)
        loss = training_loss(pkg, gt, view=cam, reg=True)
        loss.backward()
        return loss

    step(render)   # warm-up: buffers grow to their steady-state sizes
    torch.cuda.synchronize()
    torch.cuda.set_sync_debug_mode("error")
    try:
        loss = step(render)
    finally:
        torch.cuda.set_sync_debug_mode("default")
    assert math.isfinite(float(loss))

    # the reference's argument forms: the seven 0-dim CUDA scalars cost ONE transfer per settings object
    # (host_scalars), nothing after that -- forward and backward reuse the cached tuple
    rs_dev = GaussianRasterizationSettings(
        image_height=torch.tensor(275).cuda(), image_width=torch.tensor(402).cuda(),
        tanfovx=torch.tensor(math.tan(cam.FoVx * 0.5)).cuda(), tanfovy=torch.tensor(math.tan(cam.FoVy * 0.5)).cuda(),
        kernel_size=0.0, bg=bg, scale_modifier=torch.tensor(1.0).cuda(), viewmatrix=cam.world_view_transform,
        projmatrix=cam.full_proj_transform, sh_degree=torch.tensor(3).cuda(), campos=cam.camera_center,
        prefiltered=False, require_depth=True, require_coord=False, debug=False)
    act = synthetic.activate(synthetic.make_gaussians(P, device="cuda"))
    means = act["means3D"].clone().requires_grad_(True)
    m2d = torch.zeros_like(means, requires_grad=True)
    torch.cuda.set_sync_debug_mode("error")
    try:
        with pytest.raises(RuntimeError):
            host_scalars(rs_dev)            # the one transfer, detected
    finally:
        torch.cuda.set_sync_debug_mode("default")
    host_scalars(rs_dev)
    torch.cuda.set_sync_debug_mode("error")
    try:
        out = GaussianRasterizer(rs_dev)(means, m2d, act["opacities"], act["tongue_class"], shs=act["shs"],
                                         scales=act["scales"], rotations=act["rotations"])
        out[0].sum().backward()
    finally:
        torch.cuda.set_sync_debug_mode("default")
    assert float(means.grad.abs().max()) > 0
