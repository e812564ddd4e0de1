"""Host-side mirror of ``gaussian_renderer.render`` (gaussian_renderer/__init__.py:8-142) and of the
slice of ``GaussianModel`` it touches (scene/gaussian_model.py:28-72,594-603), so that the hot path
can be driven exactly as train.py drives it (train.py:178) where the reference tree is absent
(GPU box, bench, smoke). With ``e-d3dgs_b200/`` on PYTHONPATH the reference's own render() runs
unchanged on top of the drop-in modules; this file only restates its glue."""
import math
from types import SimpleNamespace

import torch
import torch.nn as nn

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
    scales = pc._scaling
    rotations = pc._rotation
    means3D_final, scales_final, rotations_final, opacity_final, shs_final, extras = pc._deformation(
        means3D, scales, rotations, opacity, time, cam_no, pc, None, shs, iter=iter, num_down_emb_c=num_down_emb_c,
        num_down_emb_f=num_down_emb_f)
    rotations_final = pc.rotation_activation(rotations_final)
    if disable_filter3D:
        scales_final = pc.scaling_activation(scales_final)
        opacity = pc.opacity_activation(opacity_final)
    else:
        scales_final, opacity = pc.apply_scaling_n_opacity_with_3D_filter(opacity=opacity_final, scales=scales_final)
    outputs = rasterizer(means3D=means3D_final, means2D=means2D, shs=shs_final if override_color is None else None,
                         colors_precomp=override_color, opacities=opacity, tongue_class=pc.tongue_class,
                         scales=scales_final, rotations=rotations_final, cov3D_precomp=None)
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


def training_loss(pkg, gt_image, reg=True):
    """Per-view loss with the reference's structure (train.py:195-271): L1 on the image plus, once
    the depth/normal regulariser is on (iteration > 15000), terms on expected/median depth, normal
    and alpha so that every rasterizer output carries an upstream gradient."""
    loss = (pkg["render"] - gt_image).abs().mean()
    if reg:
        n = pkg["normal"]
        loss = loss + 0.05 * (1.0 - n[2]).abs().mean() + 0.01 * (pkg["expected_depth"] - pkg["median_depth"]).abs().mean()
        loss = loss + 0.01 * pkg["mask"].mean()
    return loss
