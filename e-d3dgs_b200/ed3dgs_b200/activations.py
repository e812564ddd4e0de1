"""Fused per-Gaussian activations between the deformation network and the rasterizer.

One forward and one backward kernel replace the torch statements of gaussian_renderer/__init__.py:77-83
(``rotation_activation`` = F.normalize, ``scaling_activation`` = exp, ``opacity_activation`` = sigmoid,
scene/gaussian_model.py:37-45) or, with a 3D filter, ``apply_scaling_n_opacity_with_3D_filter``
(scene/gaussian_model.py:594-603). Nothing is saved for backward except the raw inputs.
"""
import torch

from . import _lib


class _Activations(torch.autograd.Function):
    @staticmethod
    def forward(ctx, scales, rotations, opacity, filter_3D):
        if not scales.is_cuda:
            raise _lib.Ed3dgsError("ed3dgs_b200 runs on CUDA tensors only (there is no CPU fallback)")
        dev = scales.device
        P = scales.shape[0]
        s = scales.detach().contiguous().float()
        q = rotations.detach().contiguous().float()
        o = opacity.detach().contiguous().float()
        f = None if filter_3D is None else filter_3D.detach().contiguous().float()
        if s.shape != (P, 3) or q.shape != (P, 4) or o.numel() != P or (f is not None and f.numel() != P):
            raise RuntimeError("activations: expected scales [P,3], rotations [P,4], opacity [P,1], filter_3D [P,1]")
        s_out, q_out, o_out = torch.empty_like(s), torch.empty_like(q), torch.empty_like(o)
        with torch.cuda.device(dev):
            _lib.check(_lib.lib().ed3dgs_activations_forward(
                P, _lib.ptr(s), _lib.ptr(q), _lib.ptr(o), _lib.ptr(f), _lib.ptr(s_out), _lib.ptr(q_out),
                _lib.ptr(o_out), _lib.stream_handle()))
        ctx.has_filter = f is not None
        ctx.save_for_backward(s, q, o, *(() if f is None else (f,)))
        ctx.set_materialize_grads(False)
        return s_out, q_out, o_out

    @staticmethod
    def backward(ctx, g_s, g_q, g_o):
        s, q, o = ctx.saved_tensors[:3]
        f = ctx.saved_tensors[3] if ctx.has_filter else None
        P = s.shape[0]
        prep = lambda g: None if g is None else g.contiguous().float()  # noqa: E731
        g_s, g_q, g_o = prep(g_s), prep(g_q), prep(g_o)
        d_s, d_q, d_o = torch.empty_like(s), torch.empty_like(q), torch.empty_like(o)
        with torch.cuda.device(s.device):
            _lib.check(_lib.lib().ed3dgs_activations_backward(
                P, _lib.ptr(s), _lib.ptr(q), _lib.ptr(o), _lib.ptr(f), _lib.ptr(g_s), _lib.ptr(g_q), _lib.ptr(g_o),
                _lib.ptr(d_s), _lib.ptr(d_q), _lib.ptr(d_o), _lib.stream_handle()))
        return d_s, d_q, d_o, None


def gaussian_activations(scales, rotations, opacity, filter_3D=None):
    """(scales, rotations, opacity) raw -> activated, as render() computes them (filter_3D=None: the
    ``disable_filter3D`` branch). Returns (scales, rotations, opacity) with the reference's shapes."""
    return _Activations.apply(scales, rotations, opacity, filter_3D)
