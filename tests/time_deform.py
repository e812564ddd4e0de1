"""Times the deformation network (forward, backward) at N Gaussians with CUDA events and prints the library's own
per-kernel event report (csrc/profiler.h). usage: python tests/time_deform.py [N] [iters] [fp32|tf32]"""
import ctypes
import sys

import torch

import conftest  # noqa: F401
from ed3dgs_b200 import _lib, synthetic
from ed3dgs_b200.renderer import GaussianScene, default_hyper


def report():
    lib = _lib.lib()
    lib.ed3dgs_profile_report.restype = ctypes.c_int
    lib.ed3dgs_profile_report.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
    buf = ctypes.create_string_buffer(1 << 16)
    lib.ed3dgs_profile_report(buf, len(buf))
    out = {}
    for ln in buf.value.decode().splitlines():
        name, ms, n = ln.rsplit(" ", 2)
        out[name] = (float(ms), int(n))
    return out


def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
    iters = int(sys.argv[2]) if len(sys.argv) > 2 else 5
    g = synthetic.make_gaussians(N, device="cuda")
    scene = GaussianScene(g, default_hyper()).cuda()
    if len(sys.argv) > 3:
        scene._deformation.precision = sys.argv[3]
    feats = scene.get_features.detach().clone().requires_grad_(True)

    def fwd():
        time_emb = torch.tensor(0.3, device="cuda").repeat(N, 1)
        return scene._deformation(scene._xyz, scene._scaling, scene._rotation, scene._opacity, time_emb, 1, scene, None,
                                  feats, iter=20000, num_down_emb_c=30, num_down_emb_f=30)

    gouts = None
    tf, tb = [], []
    for it in range(iters + 2):
        if it == 2:
            _lib.lib().ed3dgs_profile_enable(1)
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record()
        out = fwd()
        e[1].record()
        if gouts is None:
            gouts = [torch.randn_like(o) for o in out[:5]]
        torch.autograd.backward(out[:5], gouts)
        e[2].record()
        torch.cuda.synchronize()
        if it >= 2:
            tf.append(e[0].elapsed_time(e[1]))
            tb.append(e[1].elapsed_time(e[2]))
    rep = report()
    _lib.lib().ed3dgs_profile_enable(0)
    print(f"N={N}: forward {sorted(tf)[len(tf) // 2]:.3f} ms, backward {sorted(tb)[len(tb) // 2]:.3f} ms (median of {iters})")
    for k, (ms, n) in sorted(rep.items(), key=lambda kv: -kv[1][0]):
        print(f"  {k:28s} {ms / iters:8.3f} ms/iter  {n / iters:5.1f} launches/iter")


if __name__ == "__main__":
    main()
