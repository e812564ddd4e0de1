"""Generates tests/golden/filter3d_golden.pt by running the REFERENCE's own compute_3D_filter
(/root/reference/scene/gaussian_model.py:538-592) on the CPU in the dev container. The method is extracted from the
reference file with ast and compiled on its own (importing scene.gaussian_model pulls open3d / plyfile / simple_knn,
none installed); `self` is a stand-in exposing get_xyz. Run from the repo root:
    python tests/golden/make_filter3d_golden.py"""
import ast
import math
import os
import sys

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "e-d3dgs_b200"))
from ed3dgs_b200 import synthetic  # noqa: E402

REF = "/root/reference/scene/gaussian_model.py"


def reference_method():
    tree = ast.parse(open(REF).read())
    for node in ast.walk(tree):
        if isinstance(node, ast.FunctionDef) and node.name == "compute_3D_filter":
            node.decorator_list = []
            mod = ast.Module(body=[node], type_ignores=[])
            ns = {"torch": torch, "math": math}
            exec(compile(mod, REF, "exec"), ns)
            return ns["compute_3D_filter"]
    raise RuntimeError("compute_3D_filter not found")


class Holder:
    def __init__(self, xyz):
        self.get_xyz = xyz
        self.filter_3D = None


def main():
    fn = reference_method()
    recs = []
    for name, P, n_cams, W, H, spread in (("wide", 4000, 16, 802, 550, 1.0), ("narrow", 3000, 5, 200, 150, 6.0),
                                          ("one_cam", 1000, 1, 64, 48, 3.0)):
        xyz = synthetic.make_gaussians(P)["xyz"] * spread
        cams = [synthetic.make_camera(k, n_cams, W, H, 8200.0 * W / 3208.0) for k in range(n_cams)]
        h = Holder(xyz)
        with torch.no_grad():
            fn(h, cams)
        recs.append(dict(name=name, xyz=xyz, W=W, H=H, n_cams=n_cams, filter_3D=h.filter_3D.clone()))
        print(name, "filter range", float(h.filter_3D.min()), float(h.filter_3D.max()),
              "distinct", int(h.filter_3D.unique().numel()))
    torch.save(recs, os.path.join(HERE, "filter3d_golden.pt"))
    print("wrote", os.path.getsize(os.path.join(HERE, "filter3d_golden.pt")) // 1024, "KiB")


if __name__ == "__main__":
    main()
