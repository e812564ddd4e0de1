"""TEST INFRASTRUCTURE -- CPU restatement (plain torch, differentiable by autograd) of the image losses of the
reference: l1_loss / ssim (utils/loss_utils.py:24-72) and depth_double_to_normal + the normal-consistency error
(utils/graphics_utils.py:240-267, train.py:249-267). Pinned by tests/golden/loss_golden.pt, which
tests/golden/make_loss_golden.py produces by executing the reference's own functions (extracted with ast; importing
utils.loss_utils pulls lpips / torchvision). Only tests/ may import this."""
import math
from math import exp

import torch
import torch.nn.functional as F


def l1_loss(network_output, gt):
    return torch.abs((network_output - gt)).mean()                                  # loss_utils.py:28


def _window(window_size, channel):
    gauss = torch.Tensor([exp(-(x - window_size // 2) ** 2 / float(2 * 1.5 ** 2)) for x in range(window_size)])
    w1 = (gauss / gauss.sum()).unsqueeze(1)                                           # :36-38
    w2 = w1.mm(w1.t()).float().unsqueeze(0).unsqueeze(0)                              # :41-43
    return w2.expand(channel, 1, window_size, window_size).contiguous()


def ssim(img1, img2, window_size=11):
    channel = img1.size(-3)
    window = _window(window_size, channel).type_as(img1).to(img1.device)
    pad = window_size // 2
    mu1 = F.conv2d(img1, window, padding=pad, groups=channel)                        # :55-56
    mu2 = F.conv2d(img2, window, padding=pad, groups=channel)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    sigma1_sq = F.conv2d(img1 * img1, window, padding=pad, groups=channel) - mu1_sq   # :62-64
    sigma2_sq = F.conv2d(img2 * img2, window, padding=pad, groups=channel) - mu2_sq
    sigma12 = F.conv2d(img1 * img2, window, padding=pad, groups=channel) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim_map = ((2 * mu1_mu2 + C1) * (2 * sigma12 + C2)) / ((mu1_sq + mu2_sq + C1) * (sigma1_sq + sigma2_sq + C2))
    return ssim_map.mean()                                                            # :71


def depth_double_to_normal(view, depth1, depth2):
    W, H = view.image_width, view.image_height                                       # graphics_utils.py:240-254
    fx = W / (2 * math.tan(view.FoVx / 2.))
    fy = H / (2 * math.tan(view.FoVy / 2.))
    dev = depth1.device
    intrins_inv = torch.tensor([[1 / fx, 0., -W / (2 * fx)], [0., 1 / fy, -H / (2 * fy)], [0., 0., 1.0]]).float().to(dev)
    grid_x, grid_y = torch.meshgrid(torch.arange(W) + 0.5, torch.arange(H) + 0.5, indexing='xy')
    points = torch.stack([grid_x, grid_y, torch.ones_like(grid_x)], dim=0).reshape(3, -1).float().to(dev)
    rays_d = intrins_inv @ points
    points1 = (depth1.reshape(1, -1) * rays_d).reshape(3, H, W)
    points2 = (depth2.reshape(1, -1) * rays_d).reshape(3, H, W)
    pts = torch.stack([points1, points2], dim=0)                                      # :256-263
    output = torch.zeros_like(pts)
    dx = pts[..., 2:, 1:-1] - pts[..., :-2, 1:-1]
    dy = pts[..., 1:-1, 2:] - pts[..., 1:-1, :-2]
    output[..., 1:-1, 1:-1] = F.normalize(torch.cross(dx, dy, dim=1), dim=1)
    return output


def depth_normal_loss(view, expected_depth, median_depth, normal, depth_ratio=0.6):
    dn = depth_double_to_normal(view, expected_depth, median_depth)                  # train.py:256-266
    err = (1 - (normal.unsqueeze(0) * dn).sum(dim=1))
    return (1 - depth_ratio) * err[0].mean() + depth_ratio * err[1].mean()
