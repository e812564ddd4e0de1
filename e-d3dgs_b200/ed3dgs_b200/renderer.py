"""Host-side mirror of ``gaussian_renderer.render`` (gaussian_renderer/__init__.py:8-142) and of the
slice of ``GaussianModel`` it touches (scene/gaussian_model.py:28-72,594-603), so that the hot path
can be driven exactly as train.py drives it (train.py:178) where the reference tree is absent
(GPU box, bench, smoke). With ``e-d3dgs_b200/`` on PYTHONPATH the reference's own render() runs
unchanged on top of the drop-in modules; this file only restates its glue."""
import math
from types import SimpleNamespace

import torch
import torch.nn as nn

from .activations import gaussian_activations
from .deformation import deform_network
from .rasterizer import GaussianRasterizationSettings, GaussianRasterizer


def default_hyper(**kw):
    """ModelHiddenParams defaults as used by the dataset configs (arguments/__init__.py:86-109,
    arguments/nersemble/default.py:5-16)."""
    d = dict(net_width=128, defor_depth=1, temporal_embedding_dim=256, gaussian_embedding_dim=32,
             min_embeddings=30, max_embeddings=150, total_num_frames=150, c2f_temporal_iter=20000,
             zero_temporal=False, no_ds=False, no_dr=False, no_do=False, no_dc=False, no_coarse_deform=False,
             no_fine_deform=False, use_coarse_temporal_embedding=False, no_c2f_temporal_embedding=False,
             use_anneal=True, deform_from_iter=5000)
    d.update(kw)
    return SimpleNamespace(**d)


class GaussianScene(nn.Module):
    """The GaussianModel fields render() reads, as nn.Parameters with the reference's names."""

    def __init__(self, g, hyper=None, sh_degree=3):
        super().__init__()
        hyper = hyper or default_hyper()
        self.hyper = hyper
        self.active_sh_degree = sh_degree
        self.max_sh_degree = 3
        self._xyz = nn.Parameter(g["xyz"].clone())
        self._features_dc = nn.Parameter(g["features"][:, :1].clone())
        self._features_rest = nn.Parameter(g["features"][:, 1:].clone())
        self._scaling = nn.Parameter(g["scaling"].clone())
        self._rotation = nn.Parameter(g["rotation"].clone())
        self._opacity = nn.Parameter(g["opacity"].clone())
        self._embedding = nn.Parameter(g["embedding"].clone())
        self.register_buffer("tongue_class", g["tongue_class"].clone())
        self.filter_3D = None
        self.fused_activations = True   # csrc/activations.cu instead of the torch statements (same values)
        self._deformation = deform_network(W=hyper.net_width, D=hyper.defor_depth,
                                           min_embeddings=hyper.min_embeddings, max_embeddings=hyper.max_embeddings,
                                           num_frames=hyper.total_num_frames, args=hyper)
        # activations (gaussian_model.py:30-45)
        self.scaling_activation = torch.exp
        self.opacity_activation = torch.sigmoid
        self.rotation_activation = torch.nn.functional.normalize

    @property
    def get_xyz(self):
        return self._xyz

    @property
    def get_features(self):
        return torch.cat((self._features_dc, self._features_rest), dim=1)

    @property
    def get_embedding(self):
        return self._embedding

    @property
    def get_scaling(self):
        return self.scaling_activation(self._scaling)

    def get_covariance(self, scaling_modifier=1):
        """gaussian_model.py:31-35,143-144: the six upper-triangle entries of (R S)(R S)^T."""
        q = self._rotation / self._rotation.norm(dim=1, keepdim=True)
        r, x, y, z = q.unbind(1)
        R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                         2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                         2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1).view(-1, 3, 3)
        L = R @ torch.diag_embed(scaling_modifier * self.get_scaling)
        c = L @ L.transpose(1, 2)
        return torch.stack([c[:, 0, 0], c[:, 0, 1], c[:, 0, 2], c[:, 1, 1], c[:, 1, 2], c[:, 2, 2]], -1)

    @torch.no_grad()
    def compute_3D_filter(self, cameras):
        """gaussian_model.py:538-592"""
        from .filter3d import compute_3D_filter
        self.filter_3D = compute_3D_filter(self.get_xyz, cameras)

    def apply_scaling_n_opacity_with_3D_filter(self, opacity, scales):
        """gaussian_model.py:594-603"""
        opacity = self.opacity_activation(opacity)
        scales = self.scaling_activation(scales)
        scales_square = torch.square(scales)
        det1 = scales_square.prod(dim=1)
        scales_after_square = scales_square + torch.square(self.filter_3D)
        det2 = scales_after_square.prod(dim=1)
        coef = torch.sqrt(det1 / det2)
        scales = torch.sqrt(scales_after_square)
        return scales, opacity * coef[..., None]


def render(viewpoint_camera, pc, pipe, bg_color, kernel_size, scaling_modifier=1.0, require_coord=True,
           require_depth=True, override_color=None, cam_no=None, iter=None, train_coarse=False, num_down_emb_c=5,
           num_down_emb_f=5, disable_filter3D=True):
    """gaussian_renderer/__init__.py:8-142, statement for statement (the settings scalars are
    passed as python numbers instead of 0-dim CUDA tensors: same values, no hidden syncs)."""
    dev = pc.get_xyz.device
    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True, device=dev) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:  # noqa: BLE001
        pass
    means3D = pc.get_xyz
    tanfovx = math.tan(viewpoint_camera.FoVx * 0.5)
    tanfovy = math.tan(viewpoint_camera.FoVy * 0.5)
    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=tanfovx, tanfovy=tanfovy, bg=bg_color, scale_modifier=scaling_modifier,
        viewmatrix=viewpoint_camera.world_view_transform, projmatrix=viewpoint_camera.full_proj_transform,
        sh_degree=pc.active_sh_degree, campos=viewpoint_camera.camera_center, prefiltered=False,
        debug=getattr(pipe, "debug", False), kernel_size=kernel_size, require_coord=require_coord,
        require_depth=require_depth)
    # the reference repeats the scalar to [N, 1] (gaussian_renderer/__init__.py:84); the network only reads
    # element [0, 0], so one device-side element avoids a pageable H2D copy (a sync) and 4 MB of traffic
    time = torch.full((1, 1), float(viewpoint_camera.time), dtype=torch.float32, device=means3D.device)
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)
    means2D = screenspace_points
    opacity = pc._opacity
    shs = pc.get_features
    scales = None
    rotations = None
    cov3D_precomp = None
    if getattr(pipe, "compute_cov3D_python", False):
        # as in the reference (:66-67) the deformation below then receives scales = rotations = None and raises
        # TypeError on scales[:, :3] (deformation.py:107): the flag cannot be combined with the deformation network
        cov3D_precomp = pc.get_covariance(scaling_modifier)
    else:
        scales = pc._scaling
        rotations = pc._rotation
    means3D_final, scales_final, rotations_final, opacity_final, shs_final, extras = pc._deformation(
        means3D, scales, rotations, opacity, time, cam_no, pc, None, shs, iter=iter, num_down_emb_c=num_down_emb_c,
        num_down_emb_f=num_down_emb_f)
    if getattr(pc, "fused_activations", False):
        # same values as the statements below, one kernel each way (csrc/activations.cu)
        scales_final, rotations_final, opacity = gaussian_activations(
            scales_final, rotations_final, opacity_final, None if disable_filter3D else pc.filter_3D)
    else:
        rotations_final = pc.rotation_activation(rotations_final)
        if disable_filter3D:
            scales_final = pc.scaling_activation(scales_final)
            opacity = pc.opacity_activation(opacity_final)
        else:
            scales_final, opacity = pc.apply_scaling_n_opacity_with_3D_filter(opacity=opacity_final,
                                                                              scales=scales_final)
    outputs = rasterizer(means3D=means3D_final, means2D=means2D, shs=shs_final if override_color is None else None,
                         colors_precomp=override_color, opacities=opacity, tongue_class=pc.tongue_class,
                         scales=scales_final, rotations=rotations_final, cov3D_precomp=cov3D_precomp)
    assert len(outputs) == 9, "only (depth-)diff-gaussian-rasterization from RaDe-GS supported!"
    (rendered_image, radii, rendered_expected_coord, rendered_median_coord, rendered_expected_depth,
     rendered_median_depth, rendered_alpha, rendered_tongue, rendered_normal) = outputs
    return {"render": rendered_image, "mask": rendered_alpha, "expected_coord": rendered_expected_coord,
            "median_coord": rendered_median_coord, "expected_depth": rendered_expected_depth,
            "median_depth": rendered_median_depth, "viewspace_points": means2D, "visibility_filter": radii > 0,
            "radii": radii, "normal": rendered_normal, "sh_coefs_final": shs_final, "extras": extras,
            "deformed_gaussian_positions": means3D_final, "tongue_mask": rendered_tongue}


def integrate(points3D, viewpoint_camera, pc, pipe, bg_color, kernel_size, loaded_iter, scaling_modifier=1.0,
              override_color=None, num_down_emb_c=5, num_down_emb_f=5):
    """gaussian_renderer/__init__.py:551-665 (the call mesh_extract_tetrahedra.py:29,45 makes per view): deform the
    Gaussians to the camera's time, apply the 3D filter, and integrate them to the query points. pc.filter_3D must
    be set (compute_3D_filter, mesh_extract_tetrahedra.py:78)."""
    means3D = pc.get_xyz
    raster_settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=math.tan(viewpoint_camera.FoVx * 0.5), tanfovy=math.tan(viewpoint_camera.FoVy * 0.5),
        kernel_size=kernel_size, bg=bg_color, scale_modifier=scaling_modifier,
        viewmatrix=viewpoint_camera.world_view_transform, projmatrix=viewpoint_camera.full_proj_transform,
        sh_degree=pc.active_sh_degree, campos=viewpoint_camera.camera_center, prefiltered=False,
        debug=getattr(pipe, "debug", False), require_depth=True, require_coord=True)
    rasterizer = GaussianRasterizer(raster_settings=raster_settings)
    time = torch.full((1, 1), float(viewpoint_camera.time), dtype=torch.float32, device=means3D.device)
    with torch.no_grad():
        means3D_final, scales_deformed, rotations_deformed, opacity_deformed, shs_final, _extras = pc._deformation(
            means3D, pc._scaling, pc._rotation, pc._opacity, time, None, pc, None, pc.get_features, iter=loaded_iter,
            num_down_emb_c=num_down_emb_c, num_down_emb_f=num_down_emb_f)
        scales_final, opacity_final = pc.apply_scaling_n_opacity_with_3D_filter(opacity=opacity_deformed,
                                                                                scales=scales_deformed)
        rotations_final = pc.rotation_activation(rotations_deformed)
        rendered_image, alpha_integrated, color_integrated, point_coordinate, point_sdf, radii = rasterizer.integrate(
            points3D=points3D, means3D=means3D_final, means2D=None,
            shs=shs_final if override_color is None else None, colors_precomp=override_color,
            opacities=opacity_final, scales=scales_final, rotations=rotations_final, cov3D_precomp=None,
            view2gaussian_precomp=None)
    return {"render": rendered_image, "alpha_integrated": alpha_integrated, "color_integrated": color_integrated,
            "point_coordinate": point_coordinate, "point_sdf": point_sdf, "visibility_filter": radii > 0,
            "radii": radii}


def training_loss(pkg, gt_image, view=None, reg=True, lambda_dssim=0.2, lambda_depth_normal=0.05, fused=True):
    """Per-view loss with the reference's structure (train.py:195-271): Ll1 + lambda_dssim * (1 - ssim) / 2 on the
    image and, once the depth/normal regulariser is on (iteration > 15000, `reg`), lambda_depth_normal times the
    depth-normal consistency of the expected / median depth maps against the rendered normals (depth_ratio 0.6),
    plus a small alpha-mask term so that every rasterizer output carries an upstream gradient.
    fused=True uses the CUDA losses of ed3dgs_b200.loss_utils; fused=False issues the reference's torch ops
    (`view` is needed for the depth-normal term; without it that term is skipped)."""
    if fused:   # one autograd node for the whole sum (loss_utils._StepLoss)
        from . import loss_utils as L
        dn = reg and view is not None
        return L.step_loss(pkg["render"], gt_image, view=view if dn else None,
                           expected_depth=pkg["expected_depth"] if dn else None,
                           median_depth=pkg["median_depth"] if dn else None, normal=pkg["normal"] if dn else None,
                           mask=pkg["mask"] if reg else None, lambda_dssim=lambda_dssim,
                           lambda_depth_normal=lambda_depth_normal, w_mask=0.01)
    loss = _torch_losses(pkg, gt_image, view, reg, lambda_dssim, lambda_depth_normal)
    if reg:
        loss = loss + 0.01 * pkg["mask"].mean()
    return loss


def _torch_losses(pkg, gt_image, view, reg, lambda_dssim, lambda_depth_normal):
    """The same terms written with the torch ops the reference issues (utils/loss_utils.py:24-72,
    utils/graphics_utils.py:240-267); used by the GPU-vs-reference timing scripts as the reference arm's loss."""
    import torch.nn.functional as F
    img = pkg["render"]
    loss = (img - gt_image).abs().mean()
    if lambda_dssim > 0:
        C = img.shape[-3]
        g = torch.tensor([math.exp(-(x - 5) ** 2 / float(2 * 1.5 ** 2)) for x in range(11)])
        w1 = (g / g.sum()).unsqueeze(1)
        win = w1.mm(w1.t()).float()[None, None].expand(C, 1, 11, 11).contiguous().to(img.device)
        conv = lambda t: F.conv2d(t, win, padding=5, groups=C)  # noqa: E731
        mu1, mu2 = conv(img), conv(gt_image)
        s1, s2, s12 = conv(img * img) - mu1.pow(2), conv(gt_image * gt_image) - mu2.pow(2), conv(img * gt_image) - mu1 * mu2
        m = ((2 * mu1 * mu2 + 0.01 ** 2) * (2 * s12 + 0.03 ** 2)) / ((mu1.pow(2) + mu2.pow(2) + 0.01 ** 2) * (s1 + s2 + 0.03 ** 2))
        loss = loss + lambda_dssim * (1 - m.mean()) / 2
    if reg and view is not None:
        W, H = view.image_width, view.image_height
        fx, fy = W / (2 * math.tan(view.FoVx / 2.)), H / (2 * math.tan(view.FoVy / 2.))
        dev = img.device
        Kinv = torch.tensor([[1 / fx, 0., -W / (2 * fx)], [0., 1 / fy, -H / (2 * fy)], [0., 0., 1.0]]).float().to(dev)
        gx, gy = torch.meshgrid(torch.arange(W) + 0.5, torch.arange(H) + 0.5, indexing='xy')
        rays = Kinv @ torch.stack([gx, gy, torch.ones_like(gx)], dim=0).reshape(3, -1).float().to(dev)
        pts = torch.stack([(pkg["expected_depth"].reshape(1, -1) * rays).reshape(3, H, W),
                           (pkg["median_depth"].reshape(1, -1) * rays).reshape(3, H, W)], dim=0)
        out = torch.zeros_like(pts)
        dx = pts[..., 2:, 1:-1] - pts[..., :-2, 1:-1]
        dy = pts[..., 1:-1, 2:] - pts[..., 1:-1, :-2]
        out[..., 1:-1, 1:-1] = F.normalize(torch.cross(dx, dy, dim=1), dim=1)
        err = (1 - (pkg["normal"].unsqueeze(0) * out).sum(dim=1))
        loss = loss + lambda_depth_normal * (0.4 * err[0].mean() + 0.6 * err[1].mean())
    return loss
