"""Drop-in for ``simple_knn._C.distCUDA2`` (submodules/simple-knn/spatial.cu:16-25)."""
import torch

from . import _lib


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    """points: float32 CUDA [P,3] -> float32 [P], mean squared distance to the 3 nearest neighbours."""
    if not points.is_cuda:
        raise _lib.Ed3dgsError("distCUDA2 needs a CUDA tensor (there is no CPU fallback)")
    P = points.shape[0]
    pts = points.contiguous().float()
    means = torch.zeros((P,), dtype=torch.float32, device=points.device)
    if P == 0:
        return means
    lib = _lib.lib()
    nbytes = lib.ed3dgs_dist_cuda2_workspace_bytes(P)
    work = torch.empty(nbytes, dtype=torch.uint8, device=points.device)
    with torch.cuda.device(points.device):
        _lib.check(lib.ed3dgs_dist_cuda2(P, pts.data_ptr(), means.data_ptr(), work.data_ptr(), nbytes,
                                         _lib.stream_handle()))
    return means
