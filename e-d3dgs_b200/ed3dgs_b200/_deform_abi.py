"""ctypes mirror of include/ed3dgs_deform.h."""
import ctypes
from ctypes import POINTER, c_float, c_int, c_size_t, c_void_p

HEADS = 5
MAX_DEPTH = 8
HEAD_DIMS = (3, 3, 4, 1, 48)

PtrArr5 = c_void_p * HEADS
PtrArr8 = c_void_p * MAX_DEPTH
FloatArr5 = c_float * HEADS


class DeformWeights(ctypes.Structure):
    _fields_ = [("W", c_int), ("depth", c_int), ("temb_dim", c_int), ("gemb_dim", c_int),
                ("feat_w", PtrArr8), ("feat_b", PtrArr8),
                ("head_w1", PtrArr5), ("head_b1", PtrArr5), ("head_w2", PtrArr5), ("head_b2", PtrArr5),
                ("precision", c_int)]


class DeformGrads(ctypes.Structure):
    _fields_ = [("feat_w", PtrArr8), ("feat_b", PtrArr8),
                ("head_w1", PtrArr5), ("head_b1", PtrArr5), ("head_w2", PtrArr5), ("head_b2", PtrArr5)]


def declare(lib):
    P = c_void_p
    lib.ed3dgs_temporal_embed_forward.restype = c_int
    lib.ed3dgs_temporal_embed_forward.argtypes = [P, c_int, c_int, c_int, P, P, c_void_p]
    lib.ed3dgs_temporal_embed_backward.restype = c_int
    lib.ed3dgs_temporal_embed_backward.argtypes = [P, c_int, c_int, c_int, P, P, P, P, c_void_p]
    lib.ed3dgs_deform_workspace_bytes.restype = c_size_t
    lib.ed3dgs_deform_workspace_bytes.argtypes = [c_int, c_int, c_int]
    lib.ed3dgs_deform_set_emb_ready_event.restype = None
    lib.ed3dgs_deform_set_emb_ready_event.argtypes = [c_void_p]
    lib.ed3dgs_deform_set_sm_reserve_last.restype = c_int
    lib.ed3dgs_deform_set_sm_reserve_last.argtypes = [c_int]
    lib.ed3dgs_deform_set_sm_reserve.restype = c_int
    lib.ed3dgs_deform_set_sm_reserve.argtypes = [c_int]
    lib.ed3dgs_deform_forward.restype = c_int
    lib.ed3dgs_deform_forward.argtypes = [POINTER(DeformWeights), POINTER(DeformWeights), c_int, P, P, P,
                                          POINTER(PtrArr5), POINTER(PtrArr5), POINTER(PtrArr5), POINTER(FloatArr5),
                                          P, c_size_t, c_void_p]
    lib.ed3dgs_deform_backward.restype = c_int
    lib.ed3dgs_deform_backward.argtypes = [POINTER(DeformWeights), POINTER(DeformWeights), POINTER(DeformGrads),
                                           POINTER(DeformGrads), c_int, P, P, P, POINTER(PtrArr5), POINTER(PtrArr5),
                                           POINTER(FloatArr5), P, c_size_t, P, P, P, c_void_p]
