"""Diagnostic: invraycov / alpha differences between product and reference integrate."""
import sys
import torch
import conftest  # noqa: F401
import ref_binding
from ed3dgs_b200 import rasterizer as R
import test_integrate_gpu as T

P, PN, W, H = [int(x) for x in (sys.argv[1:5] if len(sys.argv) > 4 else (3000, 12000, 160, 120))]
rs, act, pts = T.make_integrate_case(P, PN, W, H)
keep = {}
orig_empty = torch.empty
def spy_empty(*a, **k):
    t = orig_empty(*a, **k)
    if k.get("dtype") == torch.uint8 and len(a) == 1 and isinstance(a[0], int) and a[0] > 0:
        keep.setdefault("bufs", []).append(t)
    return t
torch.empty = spy_empty
mine, ref = T.run_both(rs, act, pts)
torch.empty = orig_empty
work = max((b for b in keep["bufs"]), key=lambda b: 0)  # placeholder
# the workspace is the uint8 tensor whose size equals ed3dgs_integrate_workspace_bytes
from ed3dgs_b200 import _lib
nb = _lib.lib().ed3dgs_integrate_workspace_bytes(P, PN, W, H)
work = [b for b in keep["bufs"] if b.numel() == nb][0]
off = (-work.data_ptr()) % 256
inte = work[off:off + P * 32].view(torch.float32).view(P, 8)
inv_mine = inte[:, :6]
cond_mine = inte[:, 6]
vis = ref["radii"] > 0
inv_ref = ref["invraycov"]
rel = (inv_mine - inv_ref).abs() / inv_ref.abs().amax(1, keepdim=True).clamp_min(1e-30)
print("visible", int(vis.sum()), "cond mismatch", int((cond_mine[vis] != ref["condition"][:P][vis].float()).sum()))
print("invraycov rel err: max", float(rel[vis].max()), "mean", float(rel[vis].mean()))
da = (mine[1] - ref["alpha_integrated"]).abs()
print("alpha diff max", float(da.max()), "n>1e-5", int((da > 1e-5).sum()), "n>1e-6", int((da > 1e-6).sum()), "of", PN)
worst = da.argmax()
print("worst point alpha mine/ref", float(mine[1][worst]), float(ref["alpha_integrated"][worst]))

# ---- fp64 emulation of the worst points from the reference's own state ----
import numpy as np
ref_obj = ref_binding.RefRasterizer()
R_ref, out = ref_obj.integrate(rs, pts, act["means3D"], act["opacities"], shs=act["shs"], scales=act["scales"],
                           rotations=act["rotations"])
ref_obj.saved = dict(rs=rs, means3D=act["means3D"], M=16, out=out)
st = ref_obj.state()
m2d = st["means2D"].double().cpu().numpy(); co = st["conic_opacity"].double().cpu().numpy()
ts = st["ts"].double().cpu().numpy(); rp = st["ray_planes"].double().cpu().numpy()
plist = st["point_list"].cpu().numpy(); ranges = st["ranges"].cpu().numpy()
inv = out["invraycov"].double().cpu().numpy(); cond = out["condition"].cpu().numpy()
coord = out["point_coordinate"].double().cpu().numpy()
gxn = (W + 15) // 16
fx = W / (2 * rs.tanfovx); fy = H / (2 * rs.tanfovy)
view = rs.viewmatrix.double().cpu().numpy()
p3 = pts.double().cpu().numpy()
order = torch.argsort(da, descending=True)[:4].cpu().numpy()
for w in order:
    x, y = coord[w]
    px, py = int(x), int(y)
    tile = (py // 16) * gxn + (px // 16)
    a, b = ranges[tile]
    pv = p3[w] @ view[:3, :3] + view[3, :3]
    pdepth = np.sqrt((pv ** 2).sum())
    cT = np.ones(5); offx = [0, -.5, .5, -.5, .5]; offy = [0, -.5, -.5, .5, .5]
    acc, Tp = 0.0, 1.0
    notes = []
    for pos in range(a, b):
        g = plist[pos]
        used = False
        margins = []
        for k in range(5):
            dx = m2d[g, 0] - (px + .5) - offx[k]; dy = m2d[g, 1] - (py + .5) - offy[k]
            power = -0.5 * (co[g, 0] * dx * dx + co[g, 2] * dy * dy) - co[g, 1] * dx * dy
            if power > 0: continue
            al = min(0.99, co[g, 3] * np.exp(power))
            margins.append(al * 255 - 1)
            if al < 1 / 255: continue
            tt = cT[k] * (1 - al)
            margins.append(tt / 1e-4 - 1)
            if tt < 1e-4: continue
            cT[k] = tt; used = True
        if not used:
            continue
        dx = m2d[g, 0] - x; dy = m2d[g, 1] - y
        depth = ts[g] + rp[g, 0] * dx + rp[g, 1] * dy
        M = np.array([[inv[g, 0], inv[g, 1], inv[g, 2]], [inv[g, 1], inv[g, 3], inv[g, 4]], [inv[g, 2], inv[g, 4], inv[g, 5]]])
        dz = ts[g] - min(pdepth, depth)
        dl = np.array([dx, dy, dz])
        terms = np.abs(np.outer(dl, dl) * M).max()
        q = dl @ M @ dl
        al = min(0.99, co[g, 3] * np.exp(-0.5 * q))
        if al < 1 / 255:
            if al * 255 > 0.98: notes.append(("3D-threshold", g, al * 255))
            continue
        notes.append(("contrib", int(g), float(al), float(Tp), float(q), float(terms), [f"{m:.1e}" for m in margins if abs(m) < 1e-3]))
        acc += al * Tp; Tp *= 1 - al
    print("point", w, "pix", px, py, "list", b - a, "mine", float(mine[1][w]), "ref", float(out["alpha_integrated"][w]), "fp64", acc)
    for n in notes: print("   ", n)

# ---- bitwise comparison of the per-Gaussian state (integration flavour of the preprocess) ----
e = torch.empty(0)
res = R.integrate_gaussians_to_points_raw(rs.bg, pts, act["means3D"], e, act["opacities"], act["scales"], act["rotations"],
                                          1.0, e, e, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, 0.0, None,
                                          H, W, act["shs"], 3, rs.campos, False, False)
pst = ref_binding.product_state(rs, P, res[0], res[7], res[8], res[9])
vis = out["radii"] > 0
for k in ("ts", "means2D", "conic_opacity", "ray_planes", "rgb", "depths"):
    a_, b_ = pst[k][vis], st[k][vis]
    print(k, "bitwise mismatches", int((a_ != b_).sum()), "of", a_.numel(), "max abs", float((a_ - b_).abs().max()))
sd = (mine[4] - out["point_sdf"])
print("sdf mismatches", int((sd != 0).sum()), "max", float(sd.abs().max()))
