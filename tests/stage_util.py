"""Stage-wise backward parity helpers.

The backward has two stages: (A) the blend kernel, which sums per-(pixel, Gaussian) terms into
per-Gaussian intermediate gradients -- summation ORDER differs between implementations (and
between two runs of the reference: float atomics), and (B) the per-Gaussian chain rule, whose
Jacobian is badly conditioned for needle / pancake Gaussians (1/lambda_min ~ 1e8) and amplifies
stage-A rounding noise. Testing B on IDENTICAL intermediates isolates real formula differences
from amplified summation noise."""
import torch

import parity_util as pu
import ref_binding


def unpack_product_intermediates(packed, P, coord):
    """[P][16|32] packed rows (raster_state.cuh) -> reference-layout intermediate tensors."""
    V = 32 if coord else 16
    g = packed.view(torch.float32)[: P * V].view(P, V)
    z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=g.device)  # noqa: E731
    conic = z(P, 4)
    conic[:, 0], conic[:, 1], conic[:, 3] = g[:, 4], g[:, 5], g[:, 6]
    inter = dict(dL_dmeans2D=g[:, 0:3].contiguous(), dL_dopacity=g[:, 3:4].contiguous(), dL_dconic=conic,
                 dL_dts=g[:, 7:8].contiguous(), dL_dcolors=g[:, 8:11].contiguous(),
                 dL_dray_planes=g[:, 11:13].contiguous(), dL_dnormals=g[:, 13:16].contiguous(),
                 dL_dview_points=z(P, 3), dL_dcamera_planes=z(P, 6))
    if coord:
        inter["dL_dview_points"] = g[:, 16:19].contiguous()
        inter["dL_dcamera_planes"] = g[:, 19:25].contiguous()
    return inter


def product_backward_stages(rs, inp, grads_fn):
    """Runs product forward + backward keeping the packed stage-A result.
    Returns (R, out, final_grads, packed, ctx) with ctx holding what stage B needs."""
    from ed3dgs_b200 import _lib as L
    from ed3dgs_b200.rasterizer import rasterize_forward_raw
    e = torch.empty(0)
    nz = lambda t: e if t is None else t  # noqa: E731
    res = rasterize_forward_raw(rs.bg, inp["means3D"], nz(inp["colors_precomp"]), inp["opacities"],
                                inp["tongue_class"], nz(inp["scales"]), nz(inp["rotations"]), rs.scale_modifier,
                                nz(inp["cov3D_precomp"]), rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy,
                                rs.kernel_size, rs.image_height, rs.image_width, nz(inp["shs"]), rs.sh_degree,
                                rs.campos, rs.prefiltered, rs.require_coord, rs.require_depth, rs.debug)
    (R, color, coord, mcoord, alpha, tongue, normal, depth, mdepth, radii, geom, binning, img) = res
    out = dict(color=color, coord=coord, mcoord=mcoord, depth=depth, mdepth=mdepth, alpha=alpha, tongue=tongue,
               normal=normal, radii=radii)
    g = grads_fn(rs, out)
    dev = color.device
    P = inp["means3D"].shape[0]
    M = inp["shs"].shape[1] if inp["shs"] is not None else 0
    H, W = int(rs.image_height), int(rs.image_width)
    lib = L.lib()
    scratch = torch.zeros(lib.ed3dgs_rasterize_backward_scratch_bytes(P, int(rs.require_coord)), dtype=torch.uint8,
                          device=dev)
    f = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)  # noqa: E731

    def call(packed_is_input, scratch_t):
        outs = dict(dL_dmeans2D=f(P, 3), dL_dcolors=f(P, 3), dL_dopacity=f(P, 1), dL_dmeans3D=f(P, 3),
                    dL_dcov3D=f(P, 6), dL_dsh=f(P, M, 3), dL_dscales=f(P, 3), dL_drotations=f(P, 4))
        if packed_is_input:
            lib.ed3dgs_test_packed_gradients_are_input(1)
        p = L.ptr
        st = lib.ed3dgs_rasterize_backward(
            P, int(rs.sh_degree), M, R, p(rs.bg), W, H, p(inp["means3D"]), p(inp["shs"]), p(inp["colors_precomp"]),
            p(alpha), p(inp["scales"]), float(rs.scale_modifier), p(inp["rotations"]), p(inp["cov3D_precomp"]),
            p(rs.viewmatrix), p(rs.projmatrix), p(rs.campos), float(rs.tanfovx), float(rs.tanfovy),
            float(rs.kernel_size), p(radii), p(normal), p(geom), p(binning), p(img), p(g["color"]), p(g["coord"]),
            p(g["mcoord"]), p(g["depth"]), p(g["mdepth"]), p(g["alpha"]), p(g["normal"]),
            p(outs["dL_dmeans2D"]), p(outs["dL_dcolors"]), p(outs["dL_dopacity"]), p(outs["dL_dmeans3D"]),
            p(outs["dL_dcov3D"]), p(outs["dL_dsh"]), p(outs["dL_dscales"]), p(outs["dL_drotations"]),
            p(scratch_t), int(rs.require_coord), int(rs.require_depth), int(rs.debug), L.stream_handle())
        L.check(st)
        torch.cuda.synchronize()
        return outs

    final = call(False, scratch)
    return R, out, final, scratch, call


def rel(x, y):
    fin = torch.isfinite(y) & torch.isfinite(x)
    scale = float(y[fin].abs().max()) if fin.any() else 0.0
    err = float((x[fin] - y[fin]).abs().max()) if fin.any() else 0.0
    l2 = float((x[fin] - y[fin]).double().norm() / y[fin].double().norm().clamp_min(1e-30)) if fin.any() else 0.0
    return dict(max_rel=err / scale if scale > 0 else err, l2_rel=l2)


def stage_report(rs, inp, grads_fn=pu.upstream_grads):
    """Stage A: product packed intermediates vs the reference's intermediates.
    Stage B: reference per-Gaussian stage vs product per-Gaussian stage on the PRODUCT's
    intermediates (identical inputs)."""
    P = inp["means3D"].shape[0]
    R, out, final, scratch, call = product_backward_stages(rs, inp, grads_fn)
    ref = ref_binding.RefRasterizer()
    Rr, rout = ref.forward(rs, inp["means3D"], inp["opacities"], inp["tongue_class"], shs=inp["shs"],
                           colors_precomp=inp["colors_precomp"], scales=inp["scales"], rotations=inp["rotations"],
                           cov3D_precomp=inp["cov3D_precomp"])
    rgr = ref.backward(grads_fn(rs, rout))
    inter = unpack_product_intermediates(scratch, P, bool(rs.require_coord))
    rep = {}
    # the reference overwrites dL_dopacity in place in stage B, so it is not comparable here
    for k in ("dL_dmeans2D", "dL_dconic", "dL_dcolors", "dL_dts", "dL_dray_planes", "dL_dnormals",
              "dL_dview_points", "dL_dcamera_planes"):
        a = inter[k].flatten(1)
        b = rgr[k].flatten(1)
        if k == "dL_dconic":
            a, b = a[:, [0, 1, 3]], b[:, [0, 1, 3]]
        rep["A_" + k] = rel(a, b)
    # stage B on identical inputs
    inter_ref = {k: v.clone() for k, v in inter.items()}
    rb = ref_binding.ref_backward_preprocess(ref, inter_ref)
    mine = call(True, scratch)
    for k in ("dL_dopacity", "dL_dmeans3D", "dL_dcov3D", "dL_dsh", "dL_dscales", "dL_drotations"):
        rep["B_" + k] = rel(mine[k], rb[k])
    for k in final:
        rep["E2E_" + k] = rel(final[k], rgr[k])
    return rep
