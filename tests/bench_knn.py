"""Timing of the GPU K-NN (K = 20) against the CPU kd-tree the reference's o3d_knn stands for (scipy cKDTree, all
cores; Open3D is not installed). usage: bench_knn.py P"""
import json
import sys
import time

import torch

import conftest  # noqa: F401
from ed3dgs_b200 import synthetic
from ed3dgs_b200.extra_utils import knn_gpu

P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
pts = synthetic.make_gaussians(P, device="cuda")["xyz"]
knn_gpu(pts, 20)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(3):
    d, i = knn_gpu(pts, 20)
b.record()
torch.cuda.synchronize()
res = dict(P=P, K=20, ms_gpu=a.elapsed_time(b) / 3)
from scipy.spatial import cKDTree
x = pts.cpu().double().numpy()
n = min(P, 200_000)
t0 = time.perf_counter()
tree = cKDTree(x)
t1 = time.perf_counter()
tree.query(x[:n], k=21, workers=-1)
t2 = time.perf_counter()
res.update(cpu_build_s=t1 - t0, cpu_query_s_full=(t2 - t1) * P / n, cpu="scipy cKDTree, all cores, query extrapolated from %d points" % n)
print(json.dumps(res))
