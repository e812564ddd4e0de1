"""Seeded synthetic NeRSemble / N3V-shaped scenes shared by the tests,
``__graft_entry__.smoke()`` and ``bench.py``. There is no network for real datasets.

Cameras are built exactly like the reference builds them: world_view_transform =
getWorld2View2(R, T).T and full_proj_transform = world_view @ getProjectionMatrix(...).T
(scene/cameras.py:84-92, utils/graphics_utils.py:106-141).
"""
import math
from types import SimpleNamespace

import numpy as np
import torch

SEED = 6666  # the reference's own seed (train.py:531)
SH_C0 = 0.28209479177387814

# name -> (P, W, H, sh_degree, frames)   BASELINE.md section 4
CONFIGS = {
    "C1": dict(P=10_000, W=256, H=256, sh_degree=0, frames=16, focal_full=8200.0, full_w=3208),
    "C2": dict(P=300_000, W=802, H=550, sh_degree=3, frames=150, focal_full=8200.0, full_w=3208),
    "C3": dict(P=1_000_000, W=1604, H=1100, sh_degree=3, frames=150, focal_full=8200.0, full_w=3208),
    "C4": dict(P=500_000, W=1352, H=1014, sh_degree=3, frames=300, focal_full=8200.0, full_w=2704),
    "C5": dict(P=3_000_000, W=2048, H=1088, sh_degree=3, frames=50, focal_full=8200.0, full_w=4096, family="technicolor"),
}


def make_gaussians(P, seed=SEED, splat_scale=0.3, device="cpu"):
    """Raw (pre-activation) Gaussian parameters with the reference's tensor shapes
    (scene/gaussian_model.py:150-192): xyz[P,3], log-scales[P,3], quaternions[P,4],
    opacity logits[P,1], SH features[P,16,3], embedding[P,32], tongue_class[P]."""
    g = torch.Generator().manual_seed(seed)
    n_shell = int(0.8 * P)
    d = torch.randn(n_shell, 3, generator=g)
    d = d / d.norm(dim=1, keepdim=True).clamp_min(1e-8)
    radii = torch.tensor([0.09, 0.12, 0.10])
    shell = d * radii + 0.003 * torch.randn(n_shell, 3, generator=g)
    cube = (torch.rand(P - n_shell, 3, generator=g) - 0.5) * 0.4
    xyz = torch.cat([shell, cube], 0)
    log_s = math.log(0.004 * splat_scale) + 0.6 * torch.randn(P, 3, generator=g)
    log_s = log_s.clamp(math.log(2e-4 * splat_scale), math.log(0.05 * splat_scale))
    flat_axis = torch.randint(0, 3, (P,), generator=g)
    log_s[torch.arange(P), flat_axis] += math.log(0.2)
    degenerate = torch.rand(P, generator=g) < 0.01
    log_s[degenerate, flat_axis[degenerate]] = math.log(1e-5)
    rot = torch.randn(P, 4, generator=g)
    opacity = 2.0 * torch.randn(P, 1, generator=g)
    sh = torch.zeros(P, 16, 3)
    sh[:, 0] = (torch.rand(P, 3, generator=g) * 2 - 1) * 0.5 / SH_C0 * 0.5 + 0.3
    sh[:, 1:] = 0.1 * torch.randn(P, 15, 3, generator=g)
    emb = 0.1 * torch.randn(P, 32, generator=g)
    tongue = (torch.rand(P, generator=g) < 0.02).float()
    out = dict(xyz=xyz, scaling=log_s, rotation=rot, opacity=opacity, features=sh, embedding=emb,
               tongue_class=tongue)
    return {k: v.float().contiguous().to(device) for k, v in out.items()}


def _world2view(cam_pos, target=(0.0, 0.0, 0.0)):
    c = np.asarray(cam_pos, dtype=np.float64)
    z = np.asarray(target, dtype=np.float64) - c
    z /= np.linalg.norm(z)
    up = np.array([0.0, -1.0, 0.0])  # COLMAP convention: +y points down in the camera frame
    x = np.cross(-up, z)
    x /= np.linalg.norm(x)
    y = np.cross(z, x)
    Rw2c = np.stack([x, y, z], 0)
    Rt = np.zeros((4, 4))
    Rt[:3, :3] = Rw2c
    Rt[:3, 3] = -Rw2c @ c
    Rt[3, 3] = 1.0
    return np.float32(Rt)


def _projection(znear, zfar, fovX, fovY):
    tanHalfFovY = math.tan(fovY / 2)
    tanHalfFovX = math.tan(fovX / 2)
    top = tanHalfFovY * znear
    bottom = -top
    right = tanHalfFovX * znear
    left = -right
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = (zfar + znear) / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def make_camera(k, n_cams, W, H, focal, frame=0, frames=1, radius=1.0, device="cpu"):
    """Camera k of n_cams on a 120-degree arc of `radius` metres looking at the origin."""
    ang = math.radians(-60.0 + 120.0 * (k + 0.5) / n_cams)
    pos = (radius * math.sin(ang), 0.05 * math.cos(3.0 * ang), -radius * math.cos(ang))
    fovx = 2.0 * math.atan(W / (2.0 * focal))
    fovy = 2.0 * math.atan(H / (2.0 * focal))
    w2v = torch.tensor(_world2view(pos)).transpose(0, 1)
    proj = _projection(0.01, 100.0, fovx, fovy).transpose(0, 1)
    full = (w2v.unsqueeze(0).bmm(proj.unsqueeze(0))).squeeze(0)
    center = w2v.inverse()[3, :3]
    # R, T as scene/cameras.py keeps them (numpy; R is the transpose of the world-to-camera rotation, so that
    # xyz @ R + T is the camera-space position, gaussian_model.py:556-558)
    Rt = w2v.transpose(0, 1).numpy()
    return SimpleNamespace(
        R=np.ascontiguousarray(Rt[:3, :3].T), T=np.ascontiguousarray(Rt[:3, 3]), original_image=True,
        image_width=W, image_height=H, FoVx=fovx, FoVy=fovy,
        world_view_transform=w2v.contiguous().to(device), full_proj_transform=full.contiguous().to(device),
        camera_center=center.contiguous().to(device), time=float(frame) / float(max(frames, 1)), uid=k)


def config_camera(name, k=0, n_cams=16, frame=0, device="cpu"):
    c = CONFIGS[name]
    focal = c["focal_full"] * c["W"] / c["full_w"]
    return make_camera(k, n_cams, c["W"], c["H"], focal, frame=frame, frames=c["frames"], device=device)


def activate(g):
    """The activations render() applies between the MLP and the rasterizer
    (gaussian_renderer/__init__.py:77-81): normalize / exp / sigmoid."""
    rot = torch.nn.functional.normalize(g["rotation"])
    return dict(means3D=g["xyz"], scales=torch.exp(g["scaling"]), rotations=rot,
                opacities=torch.sigmoid(g["opacity"]), shs=g["features"], tongue_class=g["tongue_class"])


def raster_settings(cam, sh_degree, kernel_size=0.0, require_coord=False, require_depth=True, bg=None,
                    settings_cls=None, debug=False):
    dev = cam.world_view_transform.device
    if bg is None:
        bg = torch.zeros(3, device=dev)
    if settings_cls is None:
        from .rasterizer import GaussianRasterizationSettings as settings_cls
    return settings_cls(
        image_height=int(cam.image_height), image_width=int(cam.image_width),
        tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5), kernel_size=kernel_size,
        bg=bg, scale_modifier=1.0, viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform,
        sh_degree=sh_degree, campos=cam.camera_center, prefiltered=False, require_depth=require_depth,
        require_coord=require_coord, debug=debug)
