"""Rasterizer-only timing, product vs unmodified reference CUDA (same inputs), CUDA events.
Usage: python tests/bench_raster.py [C2|C3] [splat_scale]"""
import json
import sys

import torch

import conftest  # noqa: F401
import parity_util as pu
import ref_binding
from ed3dgs_b200 import synthetic


def timeit(fn, warm=3, iters=10):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    name = sys.argv[1] if len(sys.argv) > 1 else "C3"
    scale = float(sys.argv[2]) if len(sys.argv) > 2 else 0.3
    c = synthetic.CONFIGS[name]
    rs, inp = pu.make_case(P=c["P"], W=c["W"], H=c["H"], sh_degree=c["sh_degree"], splat_scale=scale)
    import stage_util as su
    R, out, final, scratch, call = su.product_backward_stages(rs, inp, pu.upstream_grads)
    g = pu.upstream_grads(rs, out)
    res = {"config": name, "P": c["P"], "R": R, "splat_scale": scale}
    res["mine_fwd_ms"] = timeit(lambda: pu.run_product(rs, inp, None))  # includes the state export copies
    from ed3dgs_b200.rasterizer import rasterize_forward_raw
    e = torch.empty(0)
    nz = lambda t: e if t is None else t  # noqa: E731
    fwd = lambda: rasterize_forward_raw(rs.bg, inp["means3D"], nz(inp["colors_precomp"]), inp["opacities"],  # noqa: E731
                                        inp["tongue_class"], nz(inp["scales"]), nz(inp["rotations"]),
                                        rs.scale_modifier, nz(inp["cov3D_precomp"]), rs.viewmatrix, rs.projmatrix,
                                        rs.tanfovx, rs.tanfovy, rs.kernel_size, rs.image_height, rs.image_width,
                                        nz(inp["shs"]), rs.sh_degree, rs.campos, rs.prefiltered, rs.require_coord,
                                        rs.require_depth, rs.debug)
    res["mine_fwd_ms"] = timeit(fwd)
    res["mine_bwd_ms"] = timeit(lambda: call(False, scratch))
    ref = ref_binding.RefRasterizer()

    def rfwd():
        ref.forward(rs, inp["means3D"], inp["opacities"], inp["tongue_class"], shs=inp["shs"], scales=inp["scales"],
                    rotations=inp["rotations"])
    res["ref_fwd_ms"] = timeit(rfwd)
    res["ref_bwd_ms"] = timeit(lambda: ref.backward(g))
    res["speedup_fwd"] = res["ref_fwd_ms"] / res["mine_fwd_ms"]
    res["speedup_bwd"] = res["ref_bwd_ms"] / res["mine_bwd_ms"]
    res["speedup_total"] = (res["ref_fwd_ms"] + res["ref_bwd_ms"]) / (res["mine_fwd_ms"] + res["mine_bwd_ms"])
    st = ref.state()
    res["tiles"] = pu.tile_stats(st, R, c["P"])
    print(json.dumps(res))


if __name__ == "__main__":
    main()
