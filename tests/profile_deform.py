"""Deformation network fwd+bwd once at N Gaussians (for ncu captures / launch lists)."""
import sys
from types import SimpleNamespace as NS
import torch
import conftest  # noqa: F401
from ed3dgs_b200 import synthetic
from ed3dgs_b200.renderer import GaussianScene, default_hyper

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 2
g = synthetic.make_gaussians(N, device="cuda")
scene = GaussianScene(g, default_hyper()).cuda()
for it in range(iters):
    time_emb = torch.tensor(0.3, device="cuda").repeat(N, 1)
    out = scene._deformation(scene._xyz, scene._scaling, scene._rotation, scene._opacity, time_emb, 1, scene, None,
                             scene.get_features, iter=20000, num_down_emb_c=30, num_down_emb_f=30)
    loss = sum((o * 0.37).sum() for o in out[:5])
    loss.backward()
    torch.cuda.synchronize()
print("done")
