"""Generates tests/golden/loss_golden.pt by running the REFERENCE's own loss functions on the CPU in the dev
container: l1_loss, ssim (/root/reference/utils/loss_utils.py:24-72) and depth_double_to_normal
(/root/reference/utils/graphics_utils.py:240-267) combined as train.py:195-203,249-267 combines them, with autograd
gradients. The functions are extracted from the reference files with ast (importing the modules pulls lpips /
torchvision); the two hard-coded .cuda() calls of graphics_utils.py:247,249 are neutralised for the CPU run.
Run from the repo root:  python tests/golden/make_loss_golden.py"""
import ast
import math
import os
import sys
from math import exp

import torch
import torch.nn.functional as F
from torch.autograd import Variable

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "e-d3dgs_b200"))
from ed3dgs_b200 import synthetic  # noqa: E402


def extract(path, names):
    tree = ast.parse(open(path).read())
    body = [n for n in tree.body if isinstance(n, ast.FunctionDef) and n.name in names]
    ns = {"torch": torch, "F": F, "Variable": Variable, "exp": exp, "math": math}
    exec(compile(ast.Module(body=body, type_ignores=[]), path, "exec"), ns)
    return ns


def main():
    lu = extract("/root/reference/utils/loss_utils.py", {"l1_loss", "gaussian", "create_window", "ssim", "_ssim"})
    gu = extract("/root/reference/utils/graphics_utils.py",
                 {"depths_double_to_points", "point_double_to_normal", "depth_double_to_normal"})
    torch.Tensor.cuda = lambda self, *a, **k: self   # CPU run of code with hard-coded .cuda()
    g = torch.Generator().manual_seed(123)
    recs = []
    for name, C, H, W in (("rgb_ragged", 3, 45, 70), ("rgb_small", 3, 16, 16), ("tiny", 3, 7, 9)):
        gt = torch.rand(C, H, W, generator=g)
        img = (gt + 0.2 * torch.randn(C, H, W, generator=g)).clamp(0, 1).requires_grad_(True)
        lam = 0.2
        Ll1 = lu["l1_loss"](img, gt)
        ssim_value, _ = lu["ssim"](img, gt)
        loss = Ll1 + lam * (1 - ssim_value) / 2
        (d_img,) = torch.autograd.grad(loss, img)
        view = synthetic.make_camera(1, 4, W, H, 8200.0 * W / 3208.0)
        yy, xx = torch.meshgrid(torch.arange(H).float(), torch.arange(W).float(), indexing="ij")
        de = (1.0 + 0.002 * xx + 0.003 * yy + 0.01 * torch.rand(H, W, generator=g))[None].requires_grad_(True)
        dm = (1.0 + 0.002 * xx + 0.003 * yy + 0.01 * torch.rand(H, W, generator=g))[None].requires_grad_(True)
        nr = F.normalize(torch.randn(3, H, W, generator=g), dim=0).requires_grad_(True)
        dn = gu["depth_double_to_normal"](view, de, dm)
        err = (1 - (nr.unsqueeze(0) * dn).sum(dim=1))
        dnl = (1 - 0.6) * err[0].mean() + 0.6 * err[1].mean()
        g_e, g_m, g_n = torch.autograd.grad(dnl, (de, dm, nr))
        recs.append(dict(name=name, img=img.detach(), gt=gt, lam=lam, Ll1=Ll1.detach(), ssim=ssim_value.detach(),
                         loss=loss.detach(), d_img=d_img, W=W, H=H, depth_e=de.detach(), depth_m=dm.detach(),
                         normal=nr.detach(), normals_from_depth=dn.detach(), depth_normal_loss=dnl.detach(),
                         d_depth_e=g_e, d_depth_m=g_m, d_normal=g_n))
        print(name, float(Ll1), float(ssim_value), float(dnl))
    torch.save(recs, os.path.join(HERE, "loss_golden.pt"))
    print("wrote", os.path.getsize(os.path.join(HERE, "loss_golden.pt")) // 1024, "KiB")


if __name__ == "__main__":
    main()
