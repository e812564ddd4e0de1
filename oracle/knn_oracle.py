"""TEST INFRASTRUCTURE -- CPU restatement of the reference's o3d_knn (utils/extra_utils.py:5-16).

The reference delegates the search to a third-party dependency that is not in /root/reference and not installed
here: Open3D (open3d.geometry.KDTreeFlann, which wraps nanoflann; the reference pins no version, its README asks
for `pip install open3d`). Its published behaviour for search_knn_vector_3d(p, k): the k points with the smallest
Euclidean distance to p (exact search, L2 on float64 coordinates), returned in ascending order of squared
distance; o3d_knn asks for num_knn + 1 and drops the first hit (the query point itself). This file restates that
by brute force in float64. The order of equidistant neighbours is unspecified in FLANN; here (and in the product)
ties are ordered by index. parity unpinned against Open3D itself (absent); pinned on the definition.
Only tests/ may import this."""
import numpy as np


def o3d_knn(pts, num_knn):
    pts = np.ascontiguousarray(pts, np.float64)
    P = pts.shape[0]
    sq = np.empty((P, num_knn), np.float64)
    ind = np.empty((P, num_knn), np.int64)
    for s in range(0, P, 1024):
        q = pts[s:s + 1024]
        dx = q[:, None, 0] - pts[None, :, 0]
        dy = q[:, None, 1] - pts[None, :, 1]
        dz = q[:, None, 2] - pts[None, :, 2]
        d = dx * dx + dy * dy + dz * dz
        d[np.arange(q.shape[0]), np.arange(s, s + q.shape[0])] = np.inf   # drop the query point itself
        order = np.lexsort((np.broadcast_to(np.arange(P), d.shape), d), axis=1)[:, :num_knn]
        ind[s:s + 1024] = order
        sq[s:s + 1024] = np.take_along_axis(d, order, axis=1)
    return sq, ind
