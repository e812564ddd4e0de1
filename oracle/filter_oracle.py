"""TEST INFRASTRUCTURE -- CPU restatement (torch, fp32, op for op) of GaussianModel.compute_3D_filter
(scene/gaussian_model.py:538-592). Pinned by tests/golden/filter3d_golden.pt, which
tests/golden/make_filter3d_golden.py produces by executing the reference's own method body (extracted from the
reference file with ast, no package import). Only tests/ may import this."""
import math

import torch


def compute_3D_filter(xyz, cameras):
    distance = torch.ones((xyz.shape[0]), device=xyz.device) * 100000.0        # :543
    valid_points = torch.zeros((xyz.shape[0]), device=xyz.device, dtype=torch.bool)
    focal_length = 0.
    for camera in cameras:                                                     # :548
        W, H = camera.image_width, camera.image_height
        focal_x = W / (2 * math.tan(camera.FoVx / 2.))
        focal_y = H / (2 * math.tan(camera.FoVy / 2.))
        R = torch.tensor(camera.R, device=xyz.device, dtype=torch.float32)
        T = torch.tensor(camera.T, device=xyz.device, dtype=torch.float32)
        xyz_cam = xyz @ R + T[None, :]                                         # :558
        valid_depth = xyz_cam[:, 2] > 0.2
        x, y, z = xyz_cam[:, 0], xyz_cam[:, 1], xyz_cam[:, 2]
        z = torch.clamp(z, min=0.001)
        x = x / z * focal_x + camera.image_width / 2.0
        y = y / z * focal_y + camera.image_height / 2.0
        in_screen = (x >= -0.15 * camera.image_width) & (x <= camera.image_width * 1.15) & \
                    (y >= -0.15 * camera.image_height) & (y <= 1.15 * camera.image_height)   # :573-575
        valid = valid_depth & in_screen
        distance[valid] = torch.min(distance[valid], z[valid])
        valid_points = valid_points | valid
        if focal_length < focal_x:
            focal_length = focal_x
    distance[~valid_points] = distance[valid_points].max()                     # :587
    filter_3D = distance / focal_length * (0.2 ** 0.5)
    return filter_3D[..., None]
