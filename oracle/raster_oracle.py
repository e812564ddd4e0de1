"""TEST INFRASTRUCTURE -- ctypes binding of oracle/liboracle_cpu.so (oracle_raster.cpp,
oracle_knn.cpp): the CPU restatement of the reference rasterizer and of distCUDA2.
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it."""
import ctypes
import os
import subprocess
from ctypes import c_float, c_int, c_void_p

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liboracle_cpu.so")
_lib = None


def build():
    subprocess.run(["make", "-C", HERE, "cpu"], check=True, stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        L = ctypes.CDLL(LIB)
        P = c_void_p
        L.orc_create.restype = c_void_p
        L.orc_destroy.argtypes = [c_void_p]
        L.orc_forward.restype = c_int
        L.orc_forward.argtypes = ([c_void_p, c_int, c_int, c_int, P, c_int, c_int] + [P] * 6 + [c_float] + [P] * 5
                                  + [c_float] * 3 + [P] * 9 + [c_int, c_int])
        L.orc_get_state.argtypes = [c_void_p] + [P] * 15
        L.orc_backward.argtypes = ([c_void_p, c_int, c_int, c_int, P, c_int, c_int] + [P] * 4 + [c_float] + [P] * 5
                                   + [c_float] * 3 + [P] * 2 + [P] * 7 + [P] * 8 + [c_int, c_int])
        L.orc_integrate.restype = c_int
        L.orc_integrate.argtypes = ([c_void_p, c_int, c_int, c_int, c_int, P, c_int, c_int] + [P] * 6 + [c_float]
                                    + [P] * 5 + [c_float] * 3 + [P] * 8)
        L.orc_dist_cuda2.argtypes = [c_int, P, P]
        _lib = L
    return _lib


def _p(t):
    return None if t is None or t.numel() == 0 else t.data_ptr()


def _c(t):
    return None if t is None else t.detach().cpu().float().contiguous()


class OracleRasterizer:
    def __init__(self):
        self.ctx = lib().orc_create()

    def __del__(self):
        try:
            lib().orc_destroy(self.ctx)
        except Exception:  # noqa: BLE001
            pass

    def forward(self, rs, means3D, opacities, tongue_class, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        a = dict(means3D=_c(means3D), opacities=_c(opacities), tongue=_c(tongue_class), shs=_c(shs),
                 colors=_c(colors_precomp), scales=_c(scales), rots=_c(rotations), cov=_c(cov3D_precomp),
                 bg=_c(rs.bg), view=_c(rs.viewmatrix), proj=_c(rs.projmatrix), campos=_c(rs.campos))
        P = a["means3D"].shape[0]
        H, W = int(rs.image_height), int(rs.image_width)
        z = lambda *s: torch.zeros(*s, dtype=torch.float32)  # noqa: E731
        out = dict(color=z(3, H, W), coord=z(3, H, W), mcoord=z(3, H, W), depth=z(1, H, W), mdepth=z(1, H, W),
                   alpha=z(1, H, W), tongue=z(1, H, W), normal=z(3, H, W), radii=torch.zeros(P, dtype=torch.int32))
        M = a["shs"].shape[1] if a["shs"] is not None and a["shs"].numel() else 0
        R = lib().orc_forward(self.ctx, P, int(rs.sh_degree), M, _p(a["bg"]), W, H, _p(a["means3D"]), _p(a["shs"]),
                              _p(a["colors"]), _p(a["opacities"]), _p(a["tongue"]), _p(a["scales"]),
                              float(rs.scale_modifier), _p(a["rots"]), _p(a["cov"]), _p(a["view"]), _p(a["proj"]),
                              _p(a["campos"]), float(rs.tanfovx), float(rs.tanfovy), float(rs.kernel_size),
                              _p(out["color"]), _p(out["coord"]), _p(out["mcoord"]), _p(out["depth"]),
                              _p(out["mdepth"]), _p(out["alpha"]), _p(out["tongue"]), _p(out["normal"]),
                              _p(out["radii"]), int(rs.require_coord), int(rs.require_depth))
        self.saved = dict(a=a, rs=rs, out=out, M=M, R=R, P=P)
        return R, out

    def integrate(self, rs, points3D, means3D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                  cov3D_precomp=None, kernel_size=0.0):
        """Rasterizer::integrate as GaussianRasterizer.integrate calls it (__init__.py:245-312: kernel_size 0.0).
        Returns (R, dict) with the reference's six outputs plus invraycov / condition."""
        a = dict(points=_c(points3D), means3D=_c(means3D), opacities=_c(opacities), shs=_c(shs), colors=_c(colors_precomp),
                 scales=_c(scales), rots=_c(rotations), cov=_c(cov3D_precomp), bg=_c(rs.bg), view=_c(rs.viewmatrix),
                 proj=_c(rs.projmatrix), campos=_c(rs.campos))
        P, PN = a["means3D"].shape[0], a["points"].shape[0]
        H, W = int(rs.image_height), int(rs.image_width)
        f = lambda v, *s: torch.full(s, v, dtype=torch.float32)  # noqa: E731
        out = dict(color=f(0.0, 9, H, W), alpha_integrated=f(1.0, PN), color_integrated=f(0.0, PN, 3),
                   point_coordinate=f(0.0, PN, 2), point_sdf=f(-1000.0, PN), radii=torch.zeros(P, dtype=torch.int32),
                   invraycov=f(0.0, P, 6), condition=torch.zeros(P, dtype=torch.uint8))
        M = a["shs"].shape[1] if a["shs"] is not None and a["shs"].numel() else 0
        R = 0
        if P != 0 and PN != 0:
            R = lib().orc_integrate(self.ctx, PN, P, int(rs.sh_degree), M, _p(a["bg"]), W, H, _p(a["points"]),
                                    _p(a["means3D"]), _p(a["shs"]), _p(a["colors"]), _p(a["opacities"]), _p(a["scales"]),
                                    float(rs.scale_modifier), _p(a["rots"]), _p(a["cov"]), _p(a["view"]), _p(a["proj"]),
                                    _p(a["campos"]), float(rs.tanfovx), float(rs.tanfovy), float(kernel_size),
                                    _p(out["color"]), _p(out["radii"]), _p(out["alpha_integrated"]),
                                    _p(out["color_integrated"]), _p(out["point_coordinate"]), _p(out["point_sdf"]),
                                    _p(out["invraycov"]), _p(out["condition"]))
        return R, out

    def state(self):
        s = self.saved
        P, R = s["P"], s["R"]
        rs = s["rs"]
        H, W = int(rs.image_height), int(rs.image_width)
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        f = lambda *sh: torch.zeros(*sh, dtype=torch.float32)  # noqa: E731
        i32 = lambda *sh: torch.zeros(*sh, dtype=torch.int32)  # noqa: E731
        i64 = lambda *sh: torch.zeros(*sh, dtype=torch.int64)  # noqa: E731
        st = dict(depths=f(P), means2D=f(P, 2), tiles_touched=i32(P), point_offsets=i32(P), conic_opacity=f(P, 4),
                  rgb=f(P, 3), keys_unsorted=i64(R), vals_unsorted=i32(R), keys_sorted=i64(R), point_list=i32(R),
                  ranges=i32(tiles, 2), n_contrib=i32(2, H, W), normals=f(P, 3), ray_planes=f(P, 2), ts=f(P))
        order = ["depths", "means2D", "tiles_touched", "point_offsets", "conic_opacity", "rgb", "keys_unsorted",
                 "vals_unsorted", "keys_sorted", "point_list", "ranges", "n_contrib", "normals", "ray_planes", "ts"]
        lib().orc_get_state(self.ctx, *[_p(st[k]) for k in order])
        return st

    def backward(self, grads):
        s = self.saved
        a, rs, out, M, P = s["a"], s["rs"], s["out"], s["M"], s["P"]
        H, W = int(rs.image_height), int(rs.image_width)
        z = lambda *sh: torch.zeros(*sh, dtype=torch.float32)  # noqa: E731
        g = {k: (_c(grads.get(k)) if grads.get(k) is not None else z(*out[k].shape))
             for k in ("color", "coord", "mcoord", "depth", "mdepth", "alpha", "normal")}
        res = dict(dL_dmeans2D=z(P, 3), dL_dcolors=z(P, 3), dL_dopacity=z(P, 1), dL_dmeans3D=z(P, 3),
                   dL_dcov3D=z(P, 6), dL_dsh=z(P, M, 3), dL_dscales=z(P, 3), dL_drotations=z(P, 4))
        lib().orc_backward(self.ctx, P, int(rs.sh_degree), M, _p(a["bg"]), W, H, _p(a["means3D"]), _p(a["shs"]),
                           _p(out["alpha"]), _p(a["scales"]), float(rs.scale_modifier), _p(a["rots"]), _p(a["cov"]),
                           _p(a["view"]), _p(a["proj"]), _p(a["campos"]), float(rs.tanfovx), float(rs.tanfovy),
                           float(rs.kernel_size), _p(out["radii"]), _p(out["normal"]), _p(g["color"]), _p(g["coord"]),
                           _p(g["mcoord"]), _p(g["depth"]), _p(g["mdepth"]), _p(g["alpha"]), _p(g["normal"]),
                           _p(res["dL_dmeans2D"]), _p(res["dL_dcolors"]), _p(res["dL_dopacity"]),
                           _p(res["dL_dmeans3D"]), _p(res["dL_dcov3D"]), _p(res["dL_dsh"]), _p(res["dL_dscales"]),
                           _p(res["dL_drotations"]), int(rs.require_coord), int(rs.require_depth))
        return res


def dist_cuda2(points):
    pts = _c(points)
    out = torch.zeros(pts.shape[0], dtype=torch.float32)
    lib().orc_dist_cuda2(pts.shape[0], _p(pts), _p(out))
    return out
