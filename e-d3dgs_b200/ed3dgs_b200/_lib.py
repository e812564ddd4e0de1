"""ctypes binding of libed3dgs_b200.so (C ABI declared in include/ed3dgs_b200.h).

There is no fallback: if the CUDA library is missing or a call fails, this raises.
"""
import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int, c_size_t, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libed3dgs_b200.so")

ALLOC_FN = ctypes.CFUNCTYPE(c_void_p, c_void_p, c_size_t)

_lib = None


class Ed3dgsError(RuntimeError):
    pass


def _declare(lib):
    P = c_void_p  # every device pointer travels as void*
    lib.ed3dgs_last_error.restype = c_char_p
    lib.ed3dgs_last_error.argtypes = []
    lib.ed3dgs_version.restype = c_char_p
    lib.ed3dgs_version.argtypes = []
    lib.ed3dgs_rasterize_forward.restype = c_int
    lib.ed3dgs_rasterize_forward.argtypes = (
        [ALLOC_FN, c_void_p, ALLOC_FN, c_void_p, ALLOC_FN, c_void_p, c_int, c_int, c_int, P, c_int, c_int]
        + [P] * 6 + [c_float] + [P] * 5 + [c_float, c_float, c_float, c_int] + [P] * 9
        + [c_int, c_int, c_int, c_void_p, POINTER(c_int)])
    lib.ed3dgs_rasterize_backward_scratch_bytes.restype = c_size_t
    lib.ed3dgs_rasterize_backward_scratch_bytes.argtypes = [c_int, c_int]
    lib.ed3dgs_rasterize_backward.restype = c_int
    lib.ed3dgs_rasterize_backward.argtypes = (
        [c_int, c_int, c_int, c_int, P, c_int, c_int] + [P] * 5 + [c_float] + [P] * 5
        + [c_float, c_float, c_float] + [P] * 5 + [P] * 7 + [P] * 8 + [P]
        + [c_int, c_int, c_int, c_void_p])
    lib.ed3dgs_rasterize_get_state.restype = c_int
    lib.ed3dgs_rasterize_get_state.argtypes = [c_int] * 5 + [P] * 3 + [P] * 17 + [c_void_p]
    lib.ed3dgs_integrate_workspace_bytes.restype = c_size_t
    lib.ed3dgs_integrate_workspace_bytes.argtypes = [c_int] * 4
    lib.ed3dgs_integrate.restype = c_int
    lib.ed3dgs_integrate.argtypes = (
        [ALLOC_FN, c_void_p, ALLOC_FN, c_void_p, ALLOC_FN, c_void_p, c_int, c_int, c_int, c_int, P, c_int, c_int]
        + [P] * 6 + [c_float] + [P] * 5 + [c_float, c_float, c_float, c_int] + [P] * 6
        + [P, c_size_t, c_int, c_void_p, POINTER(c_int)])
    lib.ed3dgs_mark_visible.restype = c_int
    lib.ed3dgs_mark_visible.argtypes = [c_int, P, P, P, P, c_void_p]
    lib.ed3dgs_dist_cuda2_workspace_bytes.restype = c_size_t
    lib.ed3dgs_dist_cuda2_workspace_bytes.argtypes = [c_int]
    lib.ed3dgs_dist_cuda2.restype = c_int
    lib.ed3dgs_dist_cuda2.argtypes = [c_int, P, P, P, c_size_t, c_void_p]
    lib.ed3dgs_dist_cuda2_static_max_points.restype = c_int
    lib.ed3dgs_dist_cuda2_static_max_points.argtypes = []
    lib.ed3dgs_dist_cuda2_static.restype = c_int
    lib.ed3dgs_dist_cuda2_static.argtypes = [c_int, P, P, c_void_p]
    lib.ed3dgs_compute_3d_filter.restype = c_int
    lib.ed3dgs_compute_3d_filter.argtypes = [c_int, P, c_int, P, c_float, P, P, POINTER(c_int), c_void_p]
    lib.ed3dgs_activations_forward.restype = c_int
    lib.ed3dgs_activations_forward.argtypes = [c_int] + [P] * 7 + [c_void_p]
    lib.ed3dgs_activations_backward.restype = c_int
    lib.ed3dgs_activations_backward.argtypes = [c_int] + [P] * 10 + [c_void_p]
    lib.ed3dgs_ssim_l1_forward.restype = c_int
    lib.ed3dgs_ssim_l1_forward.argtypes = [c_int, c_int, c_int, P, P, P, P, c_void_p]
    lib.ed3dgs_ssim_l1_backward.restype = c_int
    lib.ed3dgs_ssim_l1_backward.argtypes = [c_int, c_int, c_int, P, P, P, P, c_float, c_float, P, c_void_p]
    lib.ed3dgs_depth_normal_forward.restype = c_int
    lib.ed3dgs_depth_normal_forward.argtypes = [c_int, c_int, c_float, c_float, P, P, P, P, P, c_void_p]
    lib.ed3dgs_depth_normal_backward.restype = c_int
    lib.ed3dgs_depth_normal_backward.argtypes = [c_int, c_int, c_float, c_float, P, P, P, P, c_float, c_float, P, P, P,
                                                 c_void_p]
    lib.ed3dgs_knn_workspace_bytes.restype = c_size_t
    lib.ed3dgs_knn_workspace_bytes.argtypes = [c_int]
    lib.ed3dgs_knn.restype = c_int
    lib.ed3dgs_knn.argtypes = [c_int, c_int, P, P, P, P, c_size_t, c_void_p]
    try:
        from . import _deform_abi
    except ImportError:
        _deform_abi = None
    if _deform_abi is not None:
        _deform_abi.declare(lib)


def lib():
    """The loaded shared library; raises Ed3dgsError if it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise Ed3dgsError(
                f"{LIB_PATH} is missing: build it with `make -C e-d3dgs_b200` "
                "(or __graft_entry__.build()). There is no CPU or PyTorch fallback.")
        handle = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
        _declare(handle)
        _lib = handle
    return _lib


def check(status):
    if status != 0:
        msg = lib().ed3dgs_last_error().decode("utf-8", "replace")
        raise Ed3dgsError(f"ed3dgs_b200 error {status}: {msg}")


def ptr(t):
    """Device pointer of a tensor, or NULL for None / empty tensors (the reference's 'absent')."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()


def stream_handle():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
