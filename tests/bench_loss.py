"""GPU timing of the fused image losses against the reference's torch formulation (oracle/loss_oracle.py run on
the same GPU: the same ops the reference issues) at 1604x1100: forward + backward."""
import json
import os
import sys

import torch

import conftest  # noqa: F401
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
import loss_oracle  # noqa: E402
from ed3dgs_b200 import loss_utils as L, synthetic  # noqa: E402

H, W = 1100, 1604
g = torch.Generator().manual_seed(9)
gt = torch.rand(3, H, W, generator=g).cuda()
img = (gt + 0.1 * torch.randn(3, H, W, device="cuda")).clamp(0, 1).requires_grad_(True)
view = synthetic.make_camera(2, 16, W, H, 8200.0 * W / 3208.0)
d0 = (1 + 1e-3 * torch.rand(1, H, W, device="cuda")).requires_grad_(True)
d1 = (1 + 1e-3 * torch.rand(1, H, W, device="cuda")).requires_grad_(True)
n0 = torch.nn.functional.normalize(torch.randn(3, H, W, device="cuda"), dim=0).requires_grad_(True)


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def run(loss_fn, leaves):
    for t in leaves:
        t.grad = None
    loss_fn().backward()


res = {
    "photometric_fused_ms": timed(lambda: run(lambda: L.photometric_loss(img, gt, 0.2)[0], [img])),
    "photometric_torch_ms": timed(lambda: run(lambda: loss_oracle.l1_loss(img, gt) + 0.2 * (1 - loss_oracle.ssim(img, gt)) / 2, [img])),
    "depth_normal_fused_ms": timed(lambda: run(lambda: L.depth_normal_loss(view, d0, d1, n0), [d0, d1, n0])),
    "depth_normal_torch_ms": timed(lambda: run(lambda: loss_oracle.depth_normal_loss(view, d0, d1, n0), [d0, d1, n0])),
}
print(json.dumps(res))
