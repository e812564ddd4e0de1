"""TEST INFRASTRUCTURE: ctypes binding of oracle/_ref/libref_ed3dgs.so, i.e. the UNMODIFIED
reference CUDA rasterizer + simple-knn compiled for sm_100a behind oracle/ref_shim.cu.
Mirrors the reference's pybind functions (rasterize_points.cu:35-250, spatial.cu:16-25) so the
parity tests can call reference and product with the same arguments."""
import ctypes
import os
from ctypes import c_float, c_int, c_void_p

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libref_ed3dgs.so")
_lib = None


def available():
    return os.path.exists(REF_LIB)


def lib():
    global _lib
    if _lib is None:
        L = ctypes.CDLL(REF_LIB)
        P = c_void_p
        L.ref_create.restype = c_void_p
        L.ref_destroy.argtypes = [c_void_p]
        L.ref_last_error.restype = ctypes.c_char_p
        L.ref_forward.restype = c_int
        L.ref_forward.argtypes = ([c_void_p, c_int, c_int, c_int, P, c_int, c_int] + [P] * 6 + [c_float] + [P] * 5
                                  + [c_float, c_float, c_float, c_int] + [P] * 9 + [c_int, c_int, c_int])
        L.ref_backward.restype = c_int
        L.ref_backward.argtypes = ([c_void_p, c_int, c_int, c_int, c_int, P, c_int, c_int] + [P] * 5 + [c_float]
                                   + [P] * 5 + [c_float, c_float, c_float] + [P] * 2 + [P] * 7 + [P] * 14
                                   + [c_int, c_int, c_int])
        L.ref_get_state.restype = c_int
        L.ref_get_state.argtypes = [c_void_p] + [P] * 17
        L.ref_integrate.restype = c_int
        L.ref_integrate.argtypes = ([c_void_p, c_int, c_int, c_int, c_int, P, c_int, c_int] + [P] * 6 + [c_float]
                                    + [P] * 5 + [c_float, c_float, c_float] + [P] * 10)
        L.ref_dist_cuda2.argtypes = [c_int, P, P]
        L.ref_mark_visible.argtypes = [c_int, P, P, P, P]
        _lib = L
    return _lib


def _p(t):
    return None if t is None or t.numel() == 0 else t.data_ptr()


class RefRasterizer:
    """One reference rasterizer context (owns the geom/binning/image buffers)."""

    def __init__(self, sync=True):
        """sync=False (timing runs): no device synchronisation around the calls. The shim launches on the legacy
        default stream, which is also torch's current stream, so ordering holds without it; the reference's own
        blocking read-back of the instance count (rasterizer_impl.cu:359) stays."""
        self.ctx = lib().ref_create()
        self.sync = sync

    def __del__(self):
        try:
            lib().ref_destroy(self.ctx)
        except Exception:  # noqa: BLE001
            pass

    def forward(self, rs, means3D, opacities, tongue_class, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        dev = means3D.device
        P = means3D.shape[0]
        H, W = int(rs.image_height), int(rs.image_width)
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)  # noqa: E731
        out = dict(color=z(3, H, W), coord=z(3, H, W), mcoord=z(3, H, W), depth=z(1, H, W), mdepth=z(1, H, W),
                   alpha=z(1, H, W), tongue=z(1, H, W), normal=z(3, H, W),
                   radii=torch.zeros(P, dtype=torch.int32, device=dev))
        M = shs.shape[1] if shs is not None and shs.numel() else 0
        self.saved = dict(rs=rs, means3D=means3D, shs=shs, colors_precomp=colors_precomp, scales=scales,
                          rotations=rotations, cov3D_precomp=cov3D_precomp, M=M, out=out)
        if self.sync:
            torch.cuda.synchronize()
        R = lib().ref_forward(
            self.ctx, P, int(rs.sh_degree), M, _p(rs.bg), W, H, _p(means3D), _p(shs), _p(colors_precomp),
            _p(opacities), _p(tongue_class), _p(scales), float(rs.scale_modifier), _p(rotations), _p(cov3D_precomp),
            _p(rs.viewmatrix), _p(rs.projmatrix), _p(rs.campos), float(rs.tanfovx), float(rs.tanfovy),
            float(rs.kernel_size), int(rs.prefiltered), _p(out["color"]), _p(out["coord"]), _p(out["mcoord"]),
            _p(out["depth"]), _p(out["mdepth"]), _p(out["alpha"]), _p(out["tongue"]), _p(out["normal"]),
            _p(out["radii"]), int(rs.require_coord), int(rs.require_depth), int(rs.debug))
        if self.sync:
            torch.cuda.synchronize()
        if R < 0:
            raise RuntimeError(lib().ref_last_error().decode())
        self.R = R
        return R, out

    def backward(self, grads):
        """grads: dict color, coord, mcoord, depth, mdepth, alpha, normal (any missing -> zeros)."""
        s = self.saved
        rs, out = s["rs"], s["out"]
        dev = s["means3D"].device
        P = s["means3D"].shape[0]
        H, W = int(rs.image_height), int(rs.image_width)
        M = s["M"]
        z = lambda *sh: torch.zeros(*sh, dtype=torch.float32, device=dev)  # noqa: E731
        g = {k: (grads.get(k) if grads.get(k) is not None else z(*out[k].shape)).contiguous()
             for k in ("color", "coord", "mcoord", "depth", "mdepth", "alpha", "normal")}
        res = dict(dL_dmeans2D=z(P, 3), dL_dview_points=z(P, 3), dL_dconic=z(P, 2, 2), dL_dopacity=z(P, 1),
                   dL_dcolors=z(P, 3), dL_dts=z(P, 1), dL_dcamera_planes=z(P, 6), dL_dray_planes=z(P, 2),
                   dL_dnormals=z(P, 3), dL_dmeans3D=z(P, 3), dL_dcov3D=z(P, 6), dL_dsh=z(P, M, 3),
                   dL_dscales=z(P, 3), dL_drotations=z(P, 4))
        if self.sync:
            torch.cuda.synchronize()
        st = lib().ref_backward(
            self.ctx, P, int(rs.sh_degree), M, self.R, _p(rs.bg), W, H, _p(s["means3D"]), _p(s["shs"]),
            _p(s["colors_precomp"]), _p(out["alpha"]), _p(s["scales"]), float(rs.scale_modifier), _p(s["rotations"]),
            _p(s["cov3D_precomp"]), _p(rs.viewmatrix), _p(rs.projmatrix), _p(rs.campos), float(rs.tanfovx),
            float(rs.tanfovy), float(rs.kernel_size), _p(out["radii"]), _p(out["normal"]),
            _p(g["color"]), _p(g["coord"]), _p(g["mcoord"]), _p(g["depth"]), _p(g["mdepth"]), _p(g["alpha"]),
            _p(g["normal"]),
            _p(res["dL_dmeans2D"]), _p(res["dL_dview_points"]), _p(res["dL_dconic"]), _p(res["dL_dopacity"]),
            _p(res["dL_dcolors"]), _p(res["dL_dts"]), _p(res["dL_dcamera_planes"]), _p(res["dL_dray_planes"]),
            _p(res["dL_dnormals"]), _p(res["dL_dmeans3D"]), _p(res["dL_dcov3D"]), _p(res["dL_dsh"]),
            _p(res["dL_dscales"]), _p(res["dL_drotations"]),
            int(rs.require_coord), int(rs.require_depth), int(rs.debug))
        if self.sync:
            torch.cuda.synchronize()
        if st != 0:
            raise RuntimeError(lib().ref_last_error().decode())
        return res

    def integrate(self, rs, points3D, means3D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                  cov3D_precomp=None):
        """IntegrateGaussiansToPointsCUDA (rasterize_points.cu:273-392) as GaussianRasterizer.integrate calls it
        (__init__.py:245-312: kernel_size 0.0, zero subpixel_offset); outputs pre-filled like the reference."""
        dev = means3D.device
        P, PN = means3D.shape[0], points3D.shape[0]
        H, W = int(rs.image_height), int(rs.image_width)
        f = lambda v, *s: torch.full(s, v, dtype=torch.float32, device=dev)  # noqa: E731
        out = dict(color=f(0.0, 9, H, W), alpha_integrated=f(1.0, PN), color_integrated=f(0.0, PN, 3),
                   point_coordinate=f(0.0, PN, 2), point_sdf=f(-1000.0, PN),
                   radii=torch.zeros(P, dtype=torch.int32, device=dev), invraycov=f(0.0, P, 6),
                   condition=torch.zeros(max(P, PN), dtype=torch.bool, device=dev))
        accum_alpha = f(0.0, 1, H, W)
        subpixel = f(0.0, H, W, 2)
        M = shs.shape[1] if shs is not None and shs.numel() else 0
        R = 0
        if P != 0 and PN != 0:
            torch.cuda.synchronize()
            R = lib().ref_integrate(
                self.ctx, PN, P, int(rs.sh_degree), M, _p(rs.bg), W, H, _p(points3D), _p(means3D), _p(shs),
                _p(colors_precomp), _p(opacities), _p(scales), float(rs.scale_modifier), _p(rotations),
                _p(cov3D_precomp), _p(rs.viewmatrix), _p(rs.projmatrix), _p(rs.campos), float(rs.tanfovx),
                float(rs.tanfovy), 0.0, _p(subpixel), _p(out["color"]), _p(accum_alpha), _p(out["invraycov"]),
                _p(out["radii"]), _p(out["alpha_integrated"]), _p(out["color_integrated"]),
                _p(out["point_coordinate"]), _p(out["point_sdf"]), _p(out["condition"]))
            torch.cuda.synchronize()
            if R < 0:
                raise RuntimeError(lib().ref_last_error().decode())
        self.R = R
        return R, out

    def state(self):
        s = self.saved
        rs = s["rs"]
        dev = s["means3D"].device
        P, R = s["means3D"].shape[0], self.R
        H, W = int(rs.image_height), int(rs.image_width)
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        f = lambda *sh: torch.zeros(*sh, dtype=torch.float32, device=dev)  # noqa: E731
        u32 = lambda *sh: torch.zeros(*sh, dtype=torch.int32, device=dev)  # noqa: E731
        u64 = lambda *sh: torch.zeros(*sh, dtype=torch.int64, device=dev)  # noqa: E731
        st = dict(depths=f(P), means2D=f(P, 2), tiles_touched=u32(P), point_offsets=u32(P), conic_opacity=f(P, 4),
                  rgb=f(P, 3), keys_unsorted=u64(R), vals_unsorted=u32(R), keys_sorted=u64(R), point_list=u32(R),
                  ranges=u32(tiles, 2), n_contrib=u32(2, H, W), normals=f(P, 3), ray_planes=f(P, 2), ts=f(P),
                  camera_planes=f(P, 6), view_points=f(P, 3))
        order = ["depths", "means2D", "tiles_touched", "point_offsets", "conic_opacity", "rgb", "keys_unsorted",
                 "vals_unsorted", "keys_sorted", "point_list", "ranges", "n_contrib", "normals", "ray_planes", "ts",
                 "camera_planes", "view_points"]
        rc = lib().ref_get_state(self.ctx, *[_p(st[k]) for k in order])
        assert rc == 0
        return st


def ref_dist_cuda2(points):
    P = points.shape[0]
    out = torch.zeros(P, dtype=torch.float32, device=points.device)
    torch.cuda.synchronize()
    lib().ref_dist_cuda2(P, points.contiguous().data_ptr(), out.data_ptr())
    torch.cuda.synchronize()
    return out


def product_state(rs, P, R, geom, binning, img):
    """Same dict as RefRasterizer.state() from the product's opaque buffers."""
    from ed3dgs_b200 import _lib as L
    dev = geom.device
    H, W = int(rs.image_height), int(rs.image_width)
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    f = lambda *sh: torch.zeros(*sh, dtype=torch.float32, device=dev)  # noqa: E731
    u32 = lambda *sh: torch.zeros(*sh, dtype=torch.int32, device=dev)  # noqa: E731
    u64 = lambda *sh: torch.zeros(*sh, dtype=torch.int64, device=dev)  # noqa: E731
    st = dict(depths=f(P), means2D=f(P, 2), tiles_touched=u32(P), point_offsets=u32(P), conic_opacity=f(P, 4),
              rgb=f(P, 3), keys_unsorted=u64(R), vals_unsorted=u32(R), keys_sorted=u64(R), point_list=u32(R),
              ranges=u32(tiles, 2), n_contrib=u32(2, H, W), normals=f(P, 3), ray_planes=f(P, 2), ts=f(P),
              camera_planes=f(P, 6), view_points=f(P, 3))
    order = ["depths", "means2D", "tiles_touched", "point_offsets", "conic_opacity", "rgb", "keys_unsorted",
             "vals_unsorted", "keys_sorted", "point_list", "ranges", "n_contrib", "normals", "ray_planes", "ts",
             "camera_planes", "view_points"]
    L.check(L.lib().ed3dgs_rasterize_get_state(P, R, W, H, int(rs.require_coord), _p(geom), _p(binning), _p(img),
                                               *[_p(st[k]) for k in order], L.stream_handle()))
    return st


def ref_backward_preprocess(ref, inter):
    """Per-Gaussian half of the reference backward on caller-supplied intermediates (dict of
    reference-layout tensors; dL_dopacity is updated in place). Returns the final gradients."""
    s = ref.saved
    rs = s["rs"]
    dev = s["means3D"].device
    P, M = s["means3D"].shape[0], s["M"]
    H, W = int(rs.image_height), int(rs.image_width)
    z = lambda *sh: torch.zeros(*sh, dtype=torch.float32, device=dev)  # noqa: E731
    res = dict(dL_dmeans3D=z(P, 3), dL_dcov3D=z(P, 6), dL_dsh=z(P, M, 3), dL_dscales=z(P, 3), dL_drotations=z(P, 4))
    L = lib()
    L.ref_backward_preprocess.restype = c_int
    L.ref_backward_preprocess.argtypes = ([c_void_p] + [c_int] * 5 + [c_void_p] * 3 + [c_float] + [c_void_p] * 5
                                          + [c_float] * 3 + [c_void_p] * 15)
    torch.cuda.synchronize()
    rc = L.ref_backward_preprocess(
        ref.ctx, P, int(rs.sh_degree), M, W, H, _p(s["means3D"]), _p(s["shs"]), _p(s["scales"]),
        float(rs.scale_modifier), _p(s["rotations"]), _p(s["cov3D_precomp"]), _p(rs.viewmatrix), _p(rs.projmatrix),
        _p(rs.campos), float(rs.tanfovx), float(rs.tanfovy), float(rs.kernel_size), _p(s["out"]["radii"]),
        _p(inter["dL_dmeans2D"]), _p(inter["dL_dconic"]), _p(inter["dL_dview_points"]), _p(inter["dL_dopacity"]),
        _p(inter["dL_dcolors"]), _p(inter["dL_dts"]), _p(inter["dL_dcamera_planes"]), _p(inter["dL_dray_planes"]),
        _p(inter["dL_dnormals"]), _p(res["dL_dmeans3D"]), _p(res["dL_dcov3D"]), _p(res["dL_dsh"]),
        _p(res["dL_dscales"]), _p(res["dL_drotations"]))
    assert rc == 0
    res["dL_dopacity"] = inter["dL_dopacity"]
    res["dL_dmeans2D"] = inter["dL_dmeans2D"]
    res["dL_dcolors"] = inter["dL_dcolors"]
    return res
