"""Drop-in for ``GaussianModel.compute_3D_filter`` (scene/gaussian_model.py:538-592): the per-Gaussian 3D
smoothing-filter size from the set of training cameras, in one GPU pass instead of a python loop over cameras."""
import ctypes
import math

import numpy as np
import torch

from . import _lib


def camera_table(cameras, device):
    """[C,16] float32: R (as stored in camera.R, used as xyz @ R), T, focal_x, focal_y, width, height; plus the
    largest focal_x (gaussian_model.py:553-554,584-585)."""
    rows, focal_length = [], 0.0
    for cam in cameras:
        W, H = cam.image_width, cam.image_height
        fx = W / (2 * math.tan(cam.FoVx / 2.))
        fy = H / (2 * math.tan(cam.FoVy / 2.))
        R = np.asarray(cam.R, dtype=np.float32).reshape(9)
        T = np.asarray(cam.T, dtype=np.float32).reshape(3)
        rows.append(np.concatenate([R, T, np.array([fx, fy, W, H], dtype=np.float32)]))
        if focal_length < fx:
            focal_length = fx
    tab = torch.from_numpy(np.stack(rows).astype(np.float32)) if rows else torch.zeros(0, 16)
    return tab.to(device).contiguous(), float(focal_length)


@torch.no_grad()
def compute_3D_filter(xyz, cameras):
    """xyz: float32 CUDA [P,3]; cameras: objects with R, T, FoVx, FoVy, image_width, image_height (scene/cameras.py).
    Returns filter_3D [P,1] as the reference stores it."""
    if not xyz.is_cuda:
        raise _lib.Ed3dgsError("compute_3D_filter needs CUDA tensors (there is no CPU fallback)")
    dev = xyz.device
    P = xyz.shape[0]
    pts = xyz.detach().contiguous().float()
    tab, focal_length = camera_table(cameras, dev)
    out = torch.empty(P, dtype=torch.float32, device=dev)
    if P == 0:
        return out[..., None]
    scratch = torch.empty(4, dtype=torch.int32, device=dev)
    any_valid = ctypes.c_int(1)
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().ed3dgs_compute_3d_filter(P, pts.data_ptr(), tab.shape[0], tab.data_ptr(), focal_length,
                                                      out.data_ptr(), scratch.data_ptr(), ctypes.byref(any_valid),
                                                      _lib.stream_handle()))
    if not any_valid.value:
        raise RuntimeError("compute_3D_filter: no camera sees any point (max() of an empty tensor in the reference)")
    return out[..., None]
