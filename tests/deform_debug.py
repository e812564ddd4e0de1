import sys
from types import SimpleNamespace as NS
import torch
import conftest  # noqa: F401
import test_deform_gpu as T

def carve(N, W, depth):
    off = 0
    lay = {}
    def take(name, count):
        nonlocal off
        off = (off + 255) // 256 * 256
        lay[name] = (off // 4, count)
        off += count * 4
    for s in range(2):
        for l in range(depth):
            take(f"R{s}{l}", N * W)
        take(f"Rz{s}", N * 5 * W)
        take(f"G{s}", N * 80)
        take(f"c1{s}", W)
        take(f"colsum{s}", 80)
    take("dZ", N * 5 * W)
    take("dH0", N * W)
    take("dH1", N * W)
    take("Wpack", 5 * W * W)
    take("dc1", W)
    return lay

def main():
    N = int(sys.argv[1]) if len(sys.argv) > 1 else 5037
    W = 128
    rec = T.oracle_record(N=N, W=W)
    from ed3dgs_b200.deformation import deform_network
    c = rec["case"]
    args = NS(**rec["args"])
    net = deform_network(D=c["D"], W=c["W"], args=args)
    net.load_state_dict(rec["state"])
    net = net.cuda()
    leaves = {k: v.clone().cuda().requires_grad_(True) for k, v in rec["inputs"].items()}
    pc = NS(get_embedding=leaves["embedding"])
    time_emb = torch.tensor(c["t"], device="cuda").repeat(N, 1)
    out = net(leaves["point"], leaves["scales"], leaves["rotations"], leaves["opacity"], time_emb, c["cam_no"], pc,
              None, leaves["sh"], iter=c["it"])
    finals, (subs, _o) = out[:5], out[5]
    loss = 0
    for i, (a, b) in enumerate(zip(finals, subs)):
        loss = loss + (a * rec["upstream"][f"w_out{i}"].cuda()).sum() + (b * rec["upstream"][f"w_sub{i}"].cuda()).sum()
    node = finals[0].grad_fn
    work = node.saved_tensors[5]
    loss.backward(retain_graph=True)
    torch.cuda.synchronize()
    wf = work.view(torch.float32)
    lay = carve(N, W, 1)
    get = lambda n, shape: wf[lay[n][0]: lay[n][0] + lay[n][1]].view(*shape)  # noqa: E731
    dZ = get("dZ", (N, 5 * W))
    names = ["pos", "scales", "rotations", "opacity", "rgb"]
    offs = [0, 8, 16, 24, 32]
    dims = [3, 3, 4, 1, 48]
    sd = {k: v.cuda() for k, v in rec["state"].items()}
    # coarse stage was processed last -> dZ, dH hold coarse values
    G = get("G0", (N, 80))
    Rz = get("Rz0", (N, 5 * W))
    for h in range(5):
        W2 = sd[f"{names[h]}_deform_c.3.weight"].double()
        exp = (G[:, offs[h]:offs[h] + dims[h]].double() @ W2) * (Rz[:, h * W:(h + 1) * W] > 0)
        got = dZ[:, h * W:(h + 1) * W].double()
        err = (got - exp).abs()
        bad = (err > 1e-4 * exp.abs().max()).nonzero()
        print(names[h], "dZ maxerr", float(err.max()), "scale", float(exp.abs().max()), "nbad", bad.shape[0])
        if bad.shape[0]:
            rows = bad[:, 0].unique()
            cols = bad[:, 1].unique()
            print("   bad rows", rows[:20].tolist(), "... n", rows.numel(), " cols", cols[:20].tolist(), "n", cols.numel())
            r0, c0 = int(bad[0, 0]), int(bad[0, 1])
            print("   sample got", float(got[r0, c0]), "exp", float(exp[r0, c0]))
    Wpack = get("Wpack", (5 * W, W))
    R0 = get("R00", (N, W))
    dH = get("dH0", (N, W))
    dR = torch.zeros(N, W, dtype=torch.float64, device="cuda")
    for h in range(5):
        W1 = sd[f"{names[h]}_deform_c.1.weight"]
        print(names[h], "Wpack err", float((Wpack[h * W:(h + 1) * W] - W1).abs().max()))
        dR += dZ[:, h * W:(h + 1) * W].double() @ W1.double()
        gW1 = dict(net.named_parameters())[f"{names[h]}_deform_c.1.weight"].grad.double()
        eW1 = dZ[:, h * W:(h + 1) * W].double().t() @ R0.double()
        print(names[h], "dW1 vs workspace-based expectation rel", float((gW1 - eW1).abs().max() / eW1.abs().max()))
    dR = dR * (R0 > 0)
    print("dH err rel", float((dH.double() - dR).abs().max() / dR.abs().max()))
    # check G against expectation
    coef = [1.0, 1.0, 1.0, 1.0, 1.0]
    for h in range(5):
        e = (rec["upstream"][f"w_out{h}"] + rec["upstream"][f"w_sub{h}"]).reshape(N, dims[h]).cuda()
        gerr = (G[:, offs[h]:offs[h] + dims[h]] - e).abs().max()
        print(names[h], "G err", float(gerr))

main()


def wgrad_probe(N=5037, W=128):
    import ctypes
    from ed3dgs_b200 import _lib as L
    lib = L.lib()
    f = lib.ed3dgs_test_tc_gemm
    f.restype = ctypes.c_int
    vp, ci, cl = ctypes.c_void_p, ctypes.c_int, ctypes.c_long
    f.argtypes = [ci, ci, ci, ci, ci, ci, vp, cl, vp, cl, vp, cl, ci, ci, vp, ci, ci, vp]
    g = torch.Generator("cuda").manual_seed(5)
    dZ = torch.randn(N, 5 * W, device="cuda", generator=g)
    R = torch.relu(torch.randn(N, W, device="cuda", generator=g))
    for kpc in (2048, 1024, 4096, 8192):
        for h in range(5):
            C = torch.zeros(W, W, device="cuda")
            bg = torch.zeros(W, device="cuda")
            A = dZ[:, h * W:]
            st = f(1, 1, W, W, W + 16, N, A.data_ptr(), 5 * W, R.data_ptr(), W, C.data_ptr(), W, kpc, W, bg.data_ptr(), 4, 0, None)
            L.check(st)
            torch.cuda.synchronize()
            ref = dZ[:, h * W:(h + 1) * W].double().t() @ R.double()
            err = float((C.double() - ref).abs().max() / ref.abs().max())
            berr = float((bg.double() - dZ[:, h * W:(h + 1) * W].double().sum(0)).abs().max())
            print(f"kpc {kpc} head {h}: rel_err {err:.2e} bias_abs_err {berr:.2e}")


print("---- wgrad probe")
wgrad_probe()
