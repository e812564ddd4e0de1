"""Drop-in for the K-nearest-neighbour helper of ``utils/extra_utils.py`` (``o3d_knn``, :5-16) that train.py:218-229
calls for the embedding regulariser every time the number of Gaussians changed. The reference builds an Open3D
KDTreeFlann on the CPU and queries it once per point in a python loop (float64); here the search runs on the GPU
(libed3dgs_b200.so, ed3dgs_knn) and is exact: same neighbour sets, squared distances in float64."""
import numpy as np
import torch

from . import _lib


def knn_gpu(points: torch.Tensor, num_knn: int):
    """points: float32 CUDA [P,3] -> (sq_dists float64 [P,K] ascending, indices int64 [P,K]), self excluded."""
    if not points.is_cuda:
        raise _lib.Ed3dgsError("knn_gpu needs a CUDA tensor (there is no CPU fallback)")
    P = points.shape[0]
    K = int(num_knn)
    dev = points.device
    pts = points.detach().contiguous().float()
    d = torch.empty((P, K), dtype=torch.float64, device=dev)
    idx = torch.empty((P, K), dtype=torch.int32, device=dev)
    if P == 0:
        return d, idx.long()
    lib = _lib.lib()
    nbytes = lib.ed3dgs_knn_workspace_bytes(P)
    work = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.check(lib.ed3dgs_knn(P, K, pts.data_ptr(), d.data_ptr(), idx.data_ptr(), work.data_ptr(), nbytes,
                                  _lib.stream_handle()))
    return d, idx.long()


def o3d_knn(pts, num_knn):
    """Same call and return convention as the reference's o3d_knn: ``pts`` is a numpy array [P,3] (a tensor is
    accepted too), the result is ``(sq_dists, indices)`` as numpy arrays [P,num_knn]."""
    t = torch.as_tensor(np.ascontiguousarray(pts) if isinstance(pts, np.ndarray) else pts, dtype=torch.float32)
    if not t.is_cuda:
        if not torch.cuda.is_available():
            raise _lib.Ed3dgsError("o3d_knn runs on the GPU (there is no CPU fallback)")
        t = t.cuda()
    d, i = knn_gpu(t, num_knn)
    return d.cpu().numpy(), i.cpu().numpy()


def weighted_l2_loss_v2(x, y, w):
    """utils/extra_utils.py:19-20"""
    return torch.sqrt(((x - y) ** 2).sum(-1) * w + 1e-20).mean()
