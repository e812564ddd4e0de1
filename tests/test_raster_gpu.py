"""GPU parity: product CUDA path (through the C ABI) vs the unmodified reference CUDA kernels
(oracle/_ref) on identical seeded inputs. Bit-exact keys / ranges / radii, renders <= 1e-5 abs,
gradients <= 1e-4 relative."""
import pytest
import torch

import parity_util as pu
import ref_binding

pytestmark = pytest.mark.gpu

needs_ref = pytest.mark.skipif(not ref_binding.available(), reason="oracle/_ref not built")


@needs_ref
@pytest.mark.parametrize("coord,depth,ks", [(False, True, 0.0), (False, False, 0.0), (True, True, 0.1),
                                            (True, False, 0.1)])
def test_small_scene_all_variants(coord, depth, ks):
    rs, inp = pu.make_case(P=3000, W=160, H=120, require_coord=coord, require_depth=depth, kernel_size=ks)
    a = pu.run_product(rs, inp, pu.upstream_grads)
    b = pu.run_reference(rs, inp, pu.upstream_grads)
    pu.assert_report(pu.compare(rs, a, b), rs)


@needs_ref
def test_c1_config():
    rs, inp = pu.make_case(P=10_000, W=256, H=256, sh_degree=0)
    a = pu.run_product(rs, inp, pu.upstream_grads)
    b = pu.run_reference(rs, inp, pu.upstream_grads)
    pu.assert_report(pu.compare(rs, a, b), rs)


@needs_ref
def test_ragged_image_and_precomputed_inputs():
    rs, inp = pu.make_case(P=5000, W=203, H=117, precomp_color=True, precomp_cov=True, kernel_size=0.1)
    a = pu.run_product(rs, inp, pu.upstream_grads)
    b = pu.run_reference(rs, inp, pu.upstream_grads)
    pu.assert_report(pu.compare(rs, a, b), rs)


@needs_ref
def test_c2_scale_bit_exact_keys():
    rs, inp = pu.make_case(P=300_000, W=802, H=550)
    a = pu.run_product(rs, inp, pu.upstream_grads)
    b = pu.run_reference(rs, inp, pu.upstream_grads)
    rep = pu.compare(rs, a, b)
    print("C2", _grad_lines(rep))
    pu.assert_report(rep, rs, full_size=True)    # fixed bounds, as at C3 - C5; the stages are held to 1e-4 below


def _full_case(name):
    from ed3dgs_b200 import synthetic
    c = synthetic.CONFIGS[name]
    return c, pu.make_case(P=c["P"], W=c["W"], H=c["H"], sh_degree=c["sh_degree"],
                           focal=c["focal_full"] * c["W"] / c["full_w"])


def _grad_lines(rep):
    keys = ("_relerr", "_outliers", "_l2rel_trim")
    return {k[len("grad_"):]: (round(v, 7) if isinstance(v, float) else v) for k, v in rep.items()
            if k.startswith("grad_") and k.endswith(keys)}


@needs_ref
@pytest.mark.parametrize("name", ["C3", "C4", "C5"])
def test_full_size_reference_parity(name):
    """BASELINE.json's full sizes against the unmodified reference kernels: 13/14-bit tile ids, a partial last radix
    digit, R in the millions (rasterizer_impl.cu:70-111 keys, :381-386 sort, :151-173 ranges). Bit-exact bookkeeping,
    renders <= 1e-5, gradients <= 1e-4 end to end for the outputs that do not pass through the ill-conditioned
    per-Gaussian chain; for those that do, fixed bounds (parity_util.assert_report, full_size) -- their 1e-4 bound is
    asserted stage by stage in test_full_size_backward_stage_by_stage."""
    c, (rs, inp) = _full_case(name)
    a = pu.run_product(rs, inp, pu.upstream_grads)
    b = pu.run_reference(rs, inp, pu.upstream_grads)
    rep = pu.compare(rs, a, b)
    print(name, "R =", rep["num_rendered"], _grad_lines(rep))
    assert rep["num_rendered"][0] > 2 * c["P"]
    pu.assert_report(rep, rs, full_size=True)


@needs_ref
@pytest.mark.parametrize("name", ["C2", "C3", "C4", "C5"])
def test_full_size_backward_stage_by_stage(name):
    """The gradient tolerance at the full sizes, asserted where run-to-run noise cannot move the verdict: the blend stage
    (backward.cu:631-1016) against the reference's intermediates, and the per-Gaussian stage (backward.cu:145-628)
    against the reference's on identical intermediates (stage_util.stage_stats / assert_stages)."""
    import stage_util as su
    c, (rs, inp) = _full_case(name)
    rep = su.stage_stats(rs, inp)
    for st in ("A", "B"):
        print(name, st, {k: (round(v["max_rel"], 7), v["outliers"], round(v["l2_trim_k"], 7)) for k, v in rep[st].items()})
    su.assert_stages(rep)


def test_empty_and_all_culled():
    from ed3dgs_b200 import synthetic
    from ed3dgs_b200.rasterizer import GaussianRasterizer
    cam = synthetic.make_camera(0, 4, 64, 48, 200.0, device="cuda")
    rs = synthetic.raster_settings(cam, 3)
    g = synthetic.activate(synthetic.make_gaussians(64, device="cuda"))
    # everything behind the camera -> zero instances, background image
    means = g["means3D"].clone()
    means[:, 2] -= 100.0
    means.requires_grad_(True)
    m2d = torch.zeros_like(means, requires_grad=True)
    out = GaussianRasterizer(rs)(means, m2d, g["opacities"], g["tongue_class"], shs=g["shs"], scales=g["scales"],
                                 rotations=g["rotations"])
    assert len(out) == 9
    assert int(out[1].abs().sum()) == 0 and float(out[6].abs().max()) == 0.0
    out[0].sum().backward()
    assert float(means.grad.abs().max()) == 0.0
    # P == 0
    z = torch.zeros(0, 3, device="cuda")
    out = GaussianRasterizer(rs)(z, z.clone(), torch.zeros(0, 1, device="cuda"), torch.zeros(0, device="cuda"),
                                 shs=torch.zeros(0, 16, 3, device="cuda"), scales=z.clone(),
                                 rotations=torch.zeros(0, 4, device="cuda"))
    assert out[0].shape == (3, 48, 64) and float(out[0].abs().max()) == 0.0


def test_argument_errors():
    from ed3dgs_b200 import synthetic
    from ed3dgs_b200.rasterizer import GaussianRasterizer
    cam = synthetic.make_camera(0, 4, 64, 48, 200.0, device="cuda")
    rs = synthetic.raster_settings(cam, 3)
    g = synthetic.activate(synthetic.make_gaussians(8, device="cuda"))
    r = GaussianRasterizer(rs)
    with pytest.raises(Exception):
        r(g["means3D"], g["means3D"], g["opacities"], g["tongue_class"])  # neither shs nor colours
    with pytest.raises(Exception):
        r(g["means3D"], g["means3D"], g["opacities"], g["tongue_class"], shs=g["shs"], scales=g["scales"])
    with pytest.raises(RuntimeError):
        r(g["means3D"][:, :2], g["means3D"], g["opacities"], g["tongue_class"], shs=g["shs"], scales=g["scales"],
          rotations=g["rotations"])


@needs_ref
def test_dist_cuda2_bit_exact():
    from ed3dgs_b200 import distCUDA2, synthetic
    for P in (5, 1000, 1025, 50_000):
        pts = synthetic.make_gaussians(P, device="cuda")["xyz"]
        mine = distCUDA2(pts)
        ref = ref_binding.ref_dist_cuda2(pts)
        assert torch.equal(mine, ref), (P, float((mine - ref).abs().max()))


@needs_ref
@pytest.mark.parametrize("P", [1_000_000, 3_000_000])
def test_dist_cuda2_bit_exact_full_sizes(P):
    """The C3 / C5 point counts, where the Morton boxes (1024 points each) number in the thousands."""
    from ed3dgs_b200 import distCUDA2, synthetic
    pts = synthetic.make_gaussians(P, device="cuda")["xyz"]
    mine = distCUDA2(pts)
    ref = ref_binding.ref_dist_cuda2(pts)
    assert torch.equal(mine, ref), (P, int((mine != ref).sum()), float((mine - ref).abs().max()))


@needs_ref
def test_mark_visible():
    from ed3dgs_b200 import synthetic
    from ed3dgs_b200.rasterizer import GaussianRasterizer
    cam = synthetic.make_camera(1, 4, 64, 48, 200.0, device="cuda")
    rs = synthetic.raster_settings(cam, 3)
    pts = synthetic.make_gaussians(4096, device="cuda")["xyz"] * 8.0
    mine = GaussianRasterizer(rs).markVisible(pts)
    ref = torch.zeros(4096, dtype=torch.bool, device="cuda")
    ref_binding.lib().ref_mark_visible(4096, pts.data_ptr(), rs.viewmatrix.data_ptr(), rs.projmatrix.data_ptr(),
                                       ref.data_ptr())
    torch.cuda.synchronize()
    assert torch.equal(mine, ref) and 0 < int(mine.sum()) < 4096


def test_steps_free_their_buffers_without_the_cyclic_gc():
    """A training step must release its opaque geometry / binning / image buffers by reference counting alone:
    with the cyclic collector off (bench.py's timed regions, or any loop where it triggers rarely) a reference
    cycle through the resize callback kept one set of buffers per step alive (512 MB per step at C3)."""
    import gc
    from types import SimpleNamespace as NS
    from ed3dgs_b200 import synthetic
    from ed3dgs_b200.renderer import GaussianScene, default_hyper, render, training_loss
    dev = torch.device("cuda:0")
    P, W, H = 20000, 320, 240
    scene = GaussianScene(synthetic.make_gaussians(P, splat_scale=1.0, device=dev), default_hyper()).to(dev)
    cam = synthetic.make_camera(3, 8, W, H, 8200.0 * W / 3208.0 * 3.0, frame=20, frames=150, device=dev)
    bg = torch.zeros(3, device=dev)
    gt = torch.rand(3, H, W, device=dev)

    def step():
        for p in scene.parameters():
            p.grad = None
        pkg = render(cam, scene, NS(debug=False), bg, 0.0, require_coord=False, require_depth=True, cam_no=1,
                     iter=20000, num_down_emb_c=30, num_down_emb_f=30)
        training_loss(pkg, gt, view=cam).backward()

    gc.collect()
    gc.disable()
    try:
        step()
        step()
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(dev)
        for _ in range(4):
            step()
        torch.cuda.synchronize()
        after = torch.cuda.memory_allocated(dev)
    finally:
        gc.enable()
    assert after <= before, (before, after)


def test_buffer_callback_cycle_is_broken_on_cpu_objects():
    import gc
    import weakref
    from ed3dgs_b200.rasterizer import _Buffer
    gc.disable()
    try:
        b = _Buffer(torch.device("cpu"), "t")
        r = weakref.ref(b)
        b.close()
        del b
        assert r() is None
    finally:
        gc.enable()


@needs_ref
def test_needle_gaussians_survive_the_patch_cull():
    """Screen-space needles hundreds of pixels long at ~45 degrees: det = a c - b^2 of their conic cancels to a few
    percent, more than the margin of the warp-level patch cull (raster_cull.cuh), which therefore hands such splats to
    the exact per-pixel test. The render must still match the reference pixel for pixel (<= 1e-5), n_contrib exactly."""
    import math
    rs, inp = pu.make_case(P=4000, W=640, H=480)
    g = torch.Generator(device="cuda").manual_seed(21)
    n = 600
    idx = torch.randperm(4000, device="cuda", generator=g)[:n]
    scales = inp["scales"].clone()
    scales[idx] = torch.tensor([0.08, 2e-5, 2e-5], device="cuda") * (0.5 + torch.rand(n, 1, device="cuda", generator=g))
    ang = math.pi / 4 + 0.2 * (torch.rand(n, device="cuda", generator=g) - 0.5)
    rot = inp["rotations"].clone()
    rot[idx] = torch.stack([torch.cos(ang / 2), torch.zeros_like(ang), torch.zeros_like(ang), torch.sin(ang / 2)], -1)
    op = inp["opacities"].clone()
    op[idx] = 0.9
    inp = dict(inp, scales=scales.contiguous(), rotations=rot.contiguous(), opacities=op.contiguous())
    a = pu.run_product(rs, inp)
    b = pu.run_reference(rs, inp)
    rep = pu.compare(rs, a, b)
    print({k: v for k, v in rep.items() if not k.startswith("grad_")})
    pu.assert_report(rep, rs)
