"""Drop-in for the reference's ``diff_gaussian_rasterization`` Python API
(submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py:21-312):
same ``GaussianRasterizationSettings`` fields, same ``GaussianRasterizer.forward`` /
``markVisible`` signatures, same 9-tuple of outputs and same autograd contract, backed by
libed3dgs_b200.so through its C ABI. PyTorch only owns memory, streams and autograd here.
"""
import ctypes
from types import SimpleNamespace
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _lib


_HOST_FIELDS = (("image_height", int), ("image_width", int), ("tanfovx", float), ("tanfovy", float),
                ("kernel_size", float), ("scale_modifier", float), ("sh_degree", int))


def host_scalars(rs):
    """The seven settings scalars as python numbers. The reference's render() passes them as 0-dim CUDA tensors
    (gaussian_renderer/__init__.py:29-37) and its extension reads each one back on its own; here all device-resident
    ones come back in ONE transfer, once per settings object (forward, backward, integrate and markVisible share the
    cached tuple). Python numbers (what ed3dgs_b200.renderer passes) cost nothing."""
    cache = getattr(rs, "__dict__", None)
    if cache is not None and "_host" in cache:
        return cache["_host"]
    vals = [getattr(rs, name) for name, _ in _HOST_FIELDS]
    dev = [i for i, v in enumerate(vals) if isinstance(v, torch.Tensor) and v.is_cuda]
    if dev:
        got = torch.stack([vals[i].detach().reshape(()).double() for i in dev]).cpu().tolist()
        for i, g in zip(dev, got):
            vals[i] = g
    out = {}
    for (name, kind), v in zip(_HOST_FIELDS, vals):
        if isinstance(v, torch.Tensor):
            v = v.item()
        out[name] = kind(v)
    host = SimpleNamespace(**out)
    if cache is not None:
        cache["_host"] = host
    return host


def cpu_deep_copy_tuple(input_tuple):
    return tuple(item.cpu().clone() if isinstance(item, torch.Tensor) else item for item in input_tuple)


class _Buffer:
    """Growable uint8 CUDA buffer handed to the C side as a resize callback
    (replaces resizeFunctional, rasterize_points.cu:27-33)."""

    _high_water = {}   # (device, tag) -> largest bucketed request so far

    def __init__(self, device, tag=None):
        self.tensor = torch.empty(0, dtype=torch.uint8, device=device)
        self.device = device
        self.tag = tag
        self.callback = _lib.ALLOC_FN(self._resize)

    @staticmethod
    def _bucket(nbytes):
        """Requests above 1 MiB are rounded up to eight steps per octave (<= 12.5 % slack)."""
        if nbytes <= (1 << 20):
            return nbytes
        step = 1 << max(0, int(nbytes - 1).bit_length() - 4)   # 1/8 of the enclosing octave's lower bound
        return (nbytes + step - 1) // step * step

    def _size(self, nbytes):
        """The instance count, and with it the binning buffer, changes with every view. Tagged buffers are
        requested at the high-water mark of their kind, so every view asks torch's caching allocator for the same
        block sizes and gets the same blocks back; varying sizes made it split and re-malloc (cudaMalloc is a
        multi-millisecond host stall in the middle of a step)."""
        n = self._bucket(int(nbytes))
        if self.tag is None:
            return n
        key = (str(self.device), self.tag)
        n = max(n, _Buffer._high_water.get(key, 0))
        _Buffer._high_water[key] = n
        return n

    def close(self):
        """Drops the ctypes callback. It holds a bound method of this object, i.e. a reference cycle that only the
        cyclic garbage collector would break: until then the buffer (hundreds of MB) stays allocated, and a loop
        that runs with the collector off or rarely triggered grows by one set of buffers per step."""
        self.callback = None

    def _resize(self, _user, nbytes):
        try:
            if self.tensor.numel() < nbytes:
                self.tensor = torch.empty(self._size(nbytes), dtype=torch.uint8, device=self.device)
            return self.tensor.data_ptr()
        except Exception:  # noqa: BLE001 - must not propagate through the C frame
            return None


def _f32c(t, device):
    """Contiguous float32 tensor on `device`; empty tensors stay empty (-> NULL)."""
    if t is None:
        return torch.empty(0, device=device)
    if t.numel() == 0:
        return t
    if t.device != device:
        t = t.to(device)
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def rasterize_forward_raw(bg, means3D, colors_precomp, opacities, tongue_class, scales, rotations,
                          scale_modifier, cov3Ds_precomp, viewmatrix, projmatrix, tanfovx, tanfovy,
                          kernel_size, image_height, image_width, sh, degree, campos, prefiltered,
                          require_coord, require_depth, debug):
    """Same argument list and 13-tuple result as the reference's pybind ``_C.rasterize_gaussians``
    (rasterize_points.h:18-42, rasterize_points.cu:35-137)."""
    if means3D.dim() != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise _lib.Ed3dgsError("ed3dgs_b200 runs on CUDA tensors only (there is no CPU fallback)")
    dev = means3D.device
    P = means3D.shape[0]
    H, W = int(image_height), int(image_width)
    means3D = _f32c(means3D, dev)
    sh = _f32c(sh, dev)
    colors_precomp = _f32c(colors_precomp, dev)
    opacities = _f32c(opacities, dev)
    tongue_class = _f32c(tongue_class, dev)
    scales = _f32c(scales, dev)
    rotations = _f32c(rotations, dev)
    cov3Ds_precomp = _f32c(cov3Ds_precomp, dev)
    bg = _f32c(bg, dev)
    viewmatrix = _f32c(viewmatrix, dev)
    projmatrix = _f32c(projmatrix, dev)
    campos = _f32c(campos, dev)
    f32 = dict(dtype=torch.float32, device=dev)
    # One allocation for all image planes. The kernel writes every pixel of the planes it owns;
    # planes it does not own must read as zero (rasterize_points.cu:72-80).
    if P == 0:
        planes = torch.zeros((16, H, W), **f32)
    else:   # only the planes this kernel variant does not write are zero-filled
        planes = torch.empty((16, H, W), **f32)
        if not require_coord:
            planes[3:9].zero_()
        if not require_depth:
            planes[12:14].zero_()
        if not (require_coord or require_depth):
            planes[9:12].zero_()
    out_color, out_coord, out_mcoord = planes[0:3], planes[3:6], planes[6:9]
    out_normal, out_depth, out_mdepth = planes[9:12], planes[12:13], planes[13:14]
    out_alpha, out_tongue = planes[14:15], planes[15:16]
    radii = torch.zeros((P,), dtype=torch.int32, device=dev)
    geom, binning, img = _Buffer(dev, "geom"), _Buffer(dev, "binning"), _Buffer(dev, "image")
    rendered = ctypes.c_int(0)
    if P != 0:
        M = sh.shape[1] if sh.numel() != 0 else 0
        with torch.cuda.device(dev):
            status = _lib.lib().ed3dgs_rasterize_forward(
                geom.callback, None, binning.callback, None, img.callback, None, P, int(degree), int(M),
                _lib.ptr(bg), W, H, _lib.ptr(means3D), _lib.ptr(sh), _lib.ptr(colors_precomp),
                _lib.ptr(opacities), _lib.ptr(tongue_class), _lib.ptr(scales), float(scale_modifier),
                _lib.ptr(rotations), _lib.ptr(cov3Ds_precomp), _lib.ptr(viewmatrix), _lib.ptr(projmatrix),
                _lib.ptr(campos), float(tanfovx), float(tanfovy), float(kernel_size), int(bool(prefiltered)),
                _lib.ptr(out_color), _lib.ptr(out_coord), _lib.ptr(out_mcoord), _lib.ptr(out_depth),
                _lib.ptr(out_mdepth), _lib.ptr(out_alpha), _lib.ptr(out_tongue), _lib.ptr(out_normal),
                _lib.ptr(radii), int(bool(require_coord)), int(bool(require_depth)), int(bool(debug)),
                _lib.stream_handle(), ctypes.byref(rendered))
        for b in (geom, binning, img):
            b.close()
        _lib.check(status)
    return (rendered.value, out_color, out_coord, out_mcoord, out_alpha, out_tongue, out_normal, out_depth,
            out_mdepth, radii, geom.tensor, binning.tensor, img.tensor)


def rasterize_backward_raw(bg, means3D, radii, colors_precomp, scales, rotations, scale_modifier,
                           cov3Ds_precomp, viewmatrix, projmatrix, tanfovx, tanfovy, kernel_size, grad_color,
                           grad_coord, grad_mcoord, grad_depth, grad_mdepth, grad_alpha, grad_normal, normal,
                           sh, degree, campos, geomBuffer, num_rendered, binningBuffer, imgBuffer, alpha,
                           require_coord, require_depth, debug):
    """Same argument list and 8-tuple result as ``_C.rasterize_gaussians_backward``
    (rasterize_points.h:44-77, rasterize_points.cu:139-250)."""
    dev = means3D.device
    P = means3D.shape[0]
    H, W = grad_color.shape[1], grad_color.shape[2]
    M = sh.shape[1] if sh.numel() != 0 else 0
    f32 = dict(dtype=torch.float32, device=dev)
    means3D, sh, colors_precomp = _f32c(means3D, dev), _f32c(sh, dev), _f32c(colors_precomp, dev)
    scales, rotations, cov3Ds_precomp = _f32c(scales, dev), _f32c(rotations, dev), _f32c(cov3Ds_precomp, dev)
    grads_in = [_f32c(g, dev) for g in (grad_color, grad_coord, grad_mcoord, grad_depth, grad_mdepth, grad_alpha,
                                        grad_normal)]
    dL_dmeans2D = torch.empty((P, 3), **f32)
    dL_dcolors = torch.empty((P, 3), **f32)
    dL_dopacity = torch.empty((P, 1), **f32)
    dL_dmeans3D = torch.empty((P, 3), **f32)
    dL_dcov3D = torch.empty((P, 6), **f32)
    dL_dsh = torch.empty((P, M, 3), **f32) if sh.numel() else torch.zeros((P, M, 3), **f32)
    dL_dscales = torch.empty((P, 3), **f32) if scales.numel() else torch.zeros((P, 3), **f32)
    dL_drotations = torch.empty((P, 4), **f32) if rotations.numel() else torch.zeros((P, 4), **f32)
    if P != 0:
        lib = _lib.lib()
        scratch = torch.empty(lib.ed3dgs_rasterize_backward_scratch_bytes(P, int(bool(require_coord))),
                              dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            status = lib.ed3dgs_rasterize_backward(
                P, int(degree), int(M), int(num_rendered), _lib.ptr(_f32c(bg, dev)), W, H, _lib.ptr(means3D),
                _lib.ptr(sh), _lib.ptr(colors_precomp), _lib.ptr(_f32c(alpha, dev)), _lib.ptr(scales),
                float(scale_modifier), _lib.ptr(rotations), _lib.ptr(cov3Ds_precomp),
                _lib.ptr(_f32c(viewmatrix, dev)), _lib.ptr(_f32c(projmatrix, dev)), _lib.ptr(_f32c(campos, dev)),
                float(tanfovx), float(tanfovy), float(kernel_size), _lib.ptr(radii), _lib.ptr(_f32c(normal, dev)),
                _lib.ptr(geomBuffer), _lib.ptr(binningBuffer), _lib.ptr(imgBuffer),
                *[_lib.ptr(g) for g in grads_in],
                _lib.ptr(dL_dmeans2D), _lib.ptr(dL_dcolors), _lib.ptr(dL_dopacity), _lib.ptr(dL_dmeans3D),
                _lib.ptr(dL_dcov3D), _lib.ptr(dL_dsh), _lib.ptr(dL_dscales), _lib.ptr(dL_drotations),
                _lib.ptr(scratch), int(bool(require_coord)), int(bool(require_depth)), int(bool(debug)),
                _lib.stream_handle())
        _lib.check(status)
    return (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)


def mark_visible_raw(means3D, viewmatrix, projmatrix):
    """``_C.mark_visible`` (rasterize_points.cu:252-275)."""
    dev = means3D.device
    P = means3D.shape[0]
    present = torch.zeros((P,), dtype=torch.bool, device=dev)
    if P != 0:
        means3D, viewmatrix, projmatrix = _f32c(means3D, dev), _f32c(viewmatrix, dev), _f32c(projmatrix, dev)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().ed3dgs_mark_visible(P, _lib.ptr(means3D), _lib.ptr(viewmatrix),
                                                      _lib.ptr(projmatrix), present.data_ptr(),
                                                      _lib.stream_handle()))
    return present


def integrate_gaussians_to_points_raw(bg, points3D, means3D, colors_precomp, opacities, scales, rotations,
                                     scale_modifier, cov3Ds_precomp, view2gaussian_precomp, viewmatrix, projmatrix,
                                     tanfovx, tanfovy, kernel_size, subpixel_offset, image_height, image_width, sh,
                                     degree, campos, prefiltered, debug):
    """Same argument list and 10-tuple result as the reference's pybind ``_C.integrate_gaussians_to_points``
    (rasterize_points.h:84-108, rasterize_points.cu:273-392). ``view2gaussian_precomp`` and
    ``subpixel_offset`` are dead arguments in the reference (forward.cu:1162) and are ignored."""
    if means3D.dim() != 2 or means3D.shape[1] != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if points3D.dim() != 2 or points3D.shape[1] != 3:
        raise RuntimeError("points3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise _lib.Ed3dgsError("ed3dgs_b200 runs on CUDA tensors only (there is no CPU fallback)")
    dev = means3D.device
    P, PN = means3D.shape[0], points3D.shape[0]
    H, W = int(image_height), int(image_width)
    points3D, means3D = _f32c(points3D, dev), _f32c(means3D, dev)
    sh, colors_precomp, opacities = _f32c(sh, dev), _f32c(colors_precomp, dev), _f32c(opacities, dev)
    scales, rotations, cov3Ds_precomp = _f32c(scales, dev), _f32c(rotations, dev), _f32c(cov3Ds_precomp, dev)
    bg, viewmatrix, projmatrix, campos = (_f32c(bg, dev), _f32c(viewmatrix, dev), _f32c(projmatrix, dev),
                                          _f32c(campos, dev))
    f32 = dict(dtype=torch.float32, device=dev)
    geom, binning, img = _Buffer(dev), _Buffer(dev), _Buffer(dev)
    rendered = ctypes.c_int(0)
    if P != 0 and PN != 0:
        # every output is fully written by the library
        out_color = torch.empty((9, H, W), **f32)
        radii = torch.zeros((P,), dtype=torch.int32, device=dev)
        alpha_integrated = torch.empty((PN,), **f32)
        color_integrated = torch.empty((PN, 3), **f32)
        coordinate2d = torch.empty((PN, 2), **f32)
        sdf = torch.empty((PN,), **f32)
        lib = _lib.lib()
        work = torch.empty(lib.ed3dgs_integrate_workspace_bytes(P, PN, W, H), dtype=torch.uint8, device=dev)
        M = sh.shape[1] if sh.numel() != 0 else 0
        with torch.cuda.device(dev):
            status = lib.ed3dgs_integrate(
                geom.callback, None, binning.callback, None, img.callback, None, PN, P, int(degree), int(M),
                _lib.ptr(bg), W, H, _lib.ptr(points3D), _lib.ptr(means3D), _lib.ptr(sh), _lib.ptr(colors_precomp),
                _lib.ptr(opacities), _lib.ptr(scales), float(scale_modifier), _lib.ptr(rotations),
                _lib.ptr(cov3Ds_precomp), _lib.ptr(viewmatrix), _lib.ptr(projmatrix), _lib.ptr(campos),
                float(tanfovx), float(tanfovy), float(kernel_size), int(bool(prefiltered)), _lib.ptr(out_color),
                _lib.ptr(radii), _lib.ptr(alpha_integrated), _lib.ptr(color_integrated), _lib.ptr(coordinate2d),
                _lib.ptr(sdf), _lib.ptr(work), work.numel(), int(bool(debug)), _lib.stream_handle(),
                ctypes.byref(rendered))
        for b in (geom, binning, img):
            b.close()
        _lib.check(status)
    else:  # the reference launches nothing and returns its pre-filled tensors (rasterize_points.cu:312-344)
        out_color = torch.zeros((9, H, W), **f32)
        radii = torch.zeros((P,), dtype=torch.int32, device=dev)
        alpha_integrated = torch.ones((PN,), **f32)
        color_integrated = torch.zeros((PN, 3), **f32)
        coordinate2d = torch.zeros((PN, 2), **f32)
        sdf = torch.full((PN,), -1000.0, **f32)
    return (rendered.value, out_color, alpha_integrated, color_integrated, coordinate2d, sdf, radii, geom.tensor,
            binning.tensor, img.tensor)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, tongue_class, scales, rotations,
                        cov3Ds_precomp, raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, tongue_class, scales,
                                     rotations, cov3Ds_precomp, raster_settings)


class _RasterizeGaussians(torch.autograd.Function):
    """Autograd node with the reference's contract (__init__.py:46-174): gradients for
    (means3D, means2D, sh, colors_precomp, opacities, None, scales, rotations, cov3Ds_precomp, None);
    incoming grad_radii and grad_tongue are ignored."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, tongue_class, scales, rotations,
                cov3Ds_precomp, raster_settings):
        rs = raster_settings
        hs = host_scalars(rs)
        args = (rs.bg, means3D, colors_precomp, opacities, tongue_class, scales, rotations,
                hs.scale_modifier, cov3Ds_precomp, rs.viewmatrix, rs.projmatrix,
                hs.tanfovx, hs.tanfovy, hs.kernel_size,
                hs.image_height, hs.image_width, sh, hs.sh_degree,
                rs.campos, bool(rs.prefiltered), bool(rs.require_coord), bool(rs.require_depth), bool(rs.debug))
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                out = rasterize_forward_raw(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise ex
        else:
            out = rasterize_forward_raw(*args)
        (num_rendered, color, coord, mcoord, alpha, tongue, normal, depth, mdepth, radii, geomBuffer,
         binningBuffer, imgBuffer) = out
        ctx.raster_settings = raster_settings
        ctx.scalars = args
        ctx.num_rendered = num_rendered
        ctx.plane_shape = tuple(color.shape[1:])
        # outputs the loss never touched arrive as None instead of image-sized zero tensors (grad_radii, grad_tongue
        # and, for the variants that do not own them, the coord / depth planes are never read by the kernels)
        ctx.set_materialize_grads(False)
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, normal, radii, sh,
                              geomBuffer, binningBuffer, imgBuffer, alpha)
        return color, radii, coord, mcoord, depth, mdepth, alpha, tongue, normal

    @staticmethod
    def backward(ctx, grad_color, grad_radii, grad_coord, grad_mcoord, grad_depth, grad_mdepth, grad_alpha,
                 grad_tongue, grad_normal):
        rs = ctx.raster_settings
        a = ctx.scalars
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, normal, radii, sh, geomBuffer,
         binningBuffer, imgBuffer, alpha) = ctx.saved_tensors
        coord, depth = a[20], a[21]
        H, W = ctx.plane_shape

        def need(g, channels, used):
            """A missing upstream gradient: zeros where the kernel variant reads the plane, NULL where it does not."""
            if g is not None:
                return g
            if used:
                return torch.zeros((channels, H, W), dtype=torch.float32, device=means3D.device)
            return torch.empty(0, device=means3D.device)

        grad_color = need(grad_color, 3, True)
        grad_alpha = need(grad_alpha, 1, True)
        grad_coord, grad_mcoord = need(grad_coord, 3, coord), need(grad_mcoord, 3, coord)
        grad_depth, grad_mdepth = need(grad_depth, 1, depth), need(grad_mdepth, 1, depth)
        grad_normal = need(grad_normal, 3, coord or depth)
        args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, a[7], cov3Ds_precomp, rs.viewmatrix,
                rs.projmatrix, a[11], a[12], a[13], grad_color, grad_coord, grad_mcoord, grad_depth, grad_mdepth,
                grad_alpha, grad_normal, normal, sh, a[17], rs.campos, geomBuffer, ctx.num_rendered,
                binningBuffer, imgBuffer, alpha, a[20], a[21], a[22])
        if rs.debug:
            cpu_args = cpu_deep_copy_tuple(args)
            try:
                res = rasterize_backward_raw(*args)
            except Exception as ex:
                torch.save(cpu_args, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise ex
        else:
            res = rasterize_backward_raw(*args)
        (grad_means2D, grad_colors_precomp, grad_opacities, grad_means3D, grad_cov3Ds_precomp, grad_sh,
         grad_scales, grad_rotations) = res
        return (grad_means3D, grad_means2D, grad_sh, grad_colors_precomp, grad_opacities, None, grad_scales,
                grad_rotations, grad_cov3Ds_precomp, None)


class _SettingsFields(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    kernel_size: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    require_depth: bool
    require_coord: bool
    debug: bool


class GaussianRasterizationSettings(_SettingsFields):
    """The reference's NamedTuple (__init__.py:176-191), same fields in the same order. The subclass only adds an
    instance dict so host_scalars() can cache the python-number view of the scalar fields on the object."""


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return mark_visible_raw(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, tongue_class, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        rs = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = torch.Tensor([])
        shs = empty if shs is None else shs
        colors_precomp = empty if colors_precomp is None else colors_precomp
        scales = empty if scales is None else scales
        rotations = empty if rotations is None else rotations
        cov3D_precomp = empty if cov3D_precomp is None else cov3D_precomp
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, tongue_class, scales,
                                   rotations, cov3D_precomp, rs)

    def integrate(self, points3D, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                  rotations=None, cov3D_precomp=None, view2gaussian_precomp=None):
        """GOF point integration for meshing (__init__.py:245-312): returns (color[9,H,W], alpha_integrated[PN],
        color_integrated[PN,3], point_coordinate[PN,2], point_sdf[PN], radii[P]). Like the reference it is not
        differentiable and passes kernel_size 0.0 regardless of the settings (:287)."""
        rs = self.raster_settings
        if (shs is None and colors_precomp is None) or (shs is not None and colors_precomp is not None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = torch.Tensor([])
        shs = empty if shs is None else shs
        colors_precomp = empty if colors_precomp is None else colors_precomp
        scales = empty if scales is None else scales
        rotations = empty if rotations is None else rotations
        cov3D_precomp = empty if cov3D_precomp is None else cov3D_precomp
        view2gaussian_precomp = empty if view2gaussian_precomp is None else view2gaussian_precomp
        hs = host_scalars(rs)
        args = (rs.bg, points3D, means3D, colors_precomp, opacities, scales, rotations,
                hs.scale_modifier, cov3D_precomp, view2gaussian_precomp, rs.viewmatrix,
                rs.projmatrix, hs.tanfovx, hs.tanfovy, 0.0, None,
                hs.image_height, hs.image_width, shs, hs.sh_degree,
                rs.campos, bool(rs.prefiltered), bool(rs.debug))
        with torch.no_grad():
            if rs.debug:
                cpu_args = cpu_deep_copy_tuple(args)
                try:
                    out = integrate_gaussians_to_points_raw(*args)
                except Exception as ex:
                    torch.save(cpu_args, "snapshot_fw.dump")
                    print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                    raise ex
            else:
                out = integrate_gaussians_to_points_raw(*args)
        (_num_rendered, color, alpha_integrated, color_integrated, point_coordinate, point_sdf, radii, _g, _b,
         _i) = out
        return color, alpha_integrated, color_integrated, point_coordinate, point_sdf, radii
