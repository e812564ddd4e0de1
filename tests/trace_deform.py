"""Dumps the clock trace of the stage-resident backward kernel (library built with -DED3DGS_TRACE)."""
import ctypes, sys
import numpy as np
import torch
import conftest  # noqa: F401
from ed3dgs_b200 import _lib, synthetic
from ed3dgs_b200.renderer import GaussianScene, default_hyper

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
g = synthetic.make_gaussians(N, device="cuda")
scene = GaussianScene(g, default_hyper()).cuda()
feats = scene.get_features.detach().clone().requires_grad_(True)
for it in range(2):
    time_emb = torch.tensor(0.3, device="cuda").repeat(N, 1)
    out = scene._deformation(scene._xyz, scene._scaling, scene._rotation, scene._opacity, time_emb, 1, scene, None, feats,
                             iter=20000, num_down_emb_c=30, num_down_emb_f=30)
    torch.autograd.backward(out[:5], [torch.randn_like(o) for o in out[:5]])
    torch.cuda.synchronize()
lib = _lib.lib()
buf = np.zeros(8000 * 5, dtype=np.int64)
lib.ed3dgs_trace_read.argtypes = [ctypes.c_void_p, ctypes.c_int]
print("rc", lib.ed3dgs_trace_read(buf.ctypes.data, buf.size))
np.save(sys.argv[2] if len(sys.argv) > 2 else "gpurun_out/trace.npy", buf)
