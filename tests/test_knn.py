"""Embedding-regulariser K-NN (utils/extra_utils.py:5-16 o3d_knn): CPU tests of the oracle's definition, GPU
parity of ed3dgs_knn against it (exact neighbour sets, exact float64 squared distances) and size-independent
properties at 1 M points."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import knn_oracle  # noqa: E402


def test_oracle_known_answers():
    pts = np.array([[0., 0, 0], [1, 0, 0], [0, 2, 0], [0, 0, 3], [5, 5, 5], [1, 0, 0.5]])
    d, i = knn_oracle.o3d_knn(pts, 2)
    assert i[0].tolist() == [1, 5] and np.allclose(d[0], [1.0, 1.25])
    assert i[1].tolist() == [5, 0] and np.allclose(d[1], [0.25, 1.0])
    assert i[4].tolist() == [3, 2]          # (5,5,2)^2 = 54 < (5,3,5)^2 = 59
    # ties are ordered by index
    sq = np.array([[0., 0, 0], [1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0]])
    d, i = knn_oracle.o3d_knn(sq, 4)
    assert i[0].tolist() == [1, 2, 3, 4] and np.all(d[0] == 1.0)


def test_oracle_matches_scipy_kdtree():
    from scipy.spatial import cKDTree
    rng = np.random.default_rng(0)
    pts = rng.normal(size=(3000, 3)).astype(np.float32)
    d, i = knn_oracle.o3d_knn(pts, 20)
    dd, ii = cKDTree(pts.astype(np.float64)).query(pts.astype(np.float64), k=21)
    assert np.array_equal(i, ii[:, 1:])
    assert np.allclose(d, dd[:, 1:] ** 2, rtol=1e-12)


@pytest.mark.gpu
@pytest.mark.parametrize("P,K", [(21, 20), (300, 5), (5000, 20), (20_000, 20), (4097, 32)])
def test_gpu_knn_matches_oracle(P, K):
    from ed3dgs_b200 import synthetic
    from ed3dgs_b200.extra_utils import knn_gpu, o3d_knn
    pts = synthetic.make_gaussians(P, device="cuda")["xyz"]
    d, i = knn_gpu(pts, K)
    rd, ri = knn_oracle.o3d_knn(pts.cpu().numpy(), K)
    assert np.array_equal(i.cpu().numpy(), ri)
    assert np.array_equal(d.cpu().numpy(), rd), float(np.abs(d.cpu().numpy() - rd).max())
    nd, ni = o3d_knn(pts.cpu().numpy(), K)          # the reference's calling convention (numpy in, numpy out)
    assert np.array_equal(ni, ri) and nd.dtype == np.float64


@pytest.mark.gpu
def test_gpu_knn_duplicates_and_clusters():
    from ed3dgs_b200.extra_utils import knn_gpu
    g = torch.Generator().manual_seed(5)
    base = torch.randn(400, 3, generator=g)
    pts = torch.cat([base, base[:50], base[:10] + 1e-7, 100 + 1e-3 * torch.randn(200, 3, generator=g)], 0).cuda()
    d, i = knn_gpu(pts, 20)
    rd, ri = knn_oracle.o3d_knn(pts.cpu().numpy(), 20)
    assert np.array_equal(d.cpu().numpy(), rd)
    assert np.array_equal(i.cpu().numpy(), ri)


@pytest.mark.gpu
def test_gpu_knn_properties_at_1m():
    from ed3dgs_b200 import synthetic
    from ed3dgs_b200.extra_utils import knn_gpu
    P, K = 1_000_000, 20
    pts = synthetic.make_gaussians(P, device="cuda")["xyz"]
    d, i = knn_gpu(pts, K)
    assert bool((d[:, 1:] >= d[:, :-1]).all())                       # ascending
    assert bool((i != torch.arange(P, device="cuda")[:, None]).all())  # self excluded
    assert int(i.min()) >= 0 and int(i.max()) < P
    # reported distances are the distances to the reported indices
    q = pts.double()
    chk = ((q[:, None, :] - q[i[:, [0, K - 1]]]) ** 2).sum(-1)
    assert torch.allclose(chk, d[:, [0, K - 1]], rtol=1e-14, atol=0)   # torch's reduction may fuse differently
    # exactness on a random sample: nobody outside the list is closer than the K-th neighbour
    sel = torch.randint(0, P, (64,), device="cuda", generator=torch.Generator("cuda").manual_seed(1))
    full = ((q[sel, None, :] - q[None, :, :]) ** 2).sum(-1)
    full[torch.arange(64, device="cuda"), sel] = float("inf")
    kth = full.topk(K, dim=1, largest=False).values
    assert torch.allclose(kth, d[sel], rtol=1e-14, atol=0)


def test_o3d_knn_refuses_cpu():
    from ed3dgs_b200 import Ed3dgsError
    from ed3dgs_b200.extra_utils import knn_gpu
    with pytest.raises(Ed3dgsError):
        knn_gpu(torch.rand(50, 3), 5)
