"""GPU parity of the tensor-core deformation network (through the C ABI) against
 (1) the golden vectors produced by the reference's own deform_network (tests/golden), and
 (2) the CPU oracle in fp64 at a larger, ragged size.
Stated tolerance for this path (3xTF32 error-compensated tensor-core GEMMs vs fp32 SGEMM):
outputs <= 2e-5, gradients <= 1e-4, both relative to the tensor's max magnitude."""
import os
import sys
from types import SimpleNamespace as NS

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import deform_oracle  # noqa: E402

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(ROOT, "tests", "golden", "deform_golden.pt")
OUT_TOL, GRAD_TOL = 2e-5, 1e-4


def _rel(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    scale = float(b.abs().max())
    return float((a - b).abs().max()) / (scale if scale > 0 else 1.0)


def run_product(rec, device="cuda", precision=None):
    from ed3dgs_b200.deformation import deform_network
    c = rec["case"]
    args = NS(**rec["args"])
    net = deform_network(D=c["D"], W=c["W"], min_embeddings=30, max_embeddings=150, num_frames=300, args=args)
    if precision is not None:
        net.precision = precision
    net.load_state_dict(rec["state"])
    net = net.to(device)
    leaves = {k: v.clone().to(device).requires_grad_(True) for k, v in rec["inputs"].items()}
    N = leaves["point"].shape[0]
    pc = NS(get_embedding=leaves["embedding"])
    time_emb = torch.tensor(c["t"], device=device).repeat(N, 1)
    out = net(leaves["point"], leaves["scales"], leaves["rotations"], leaves["opacity"], time_emb, c["cam_no"], pc,
              None, leaves["sh"], iter=c["it"], num_down_emb_c=30, num_down_emb_f=30)
    finals, (subs, _o) = out[:5], out[5]
    loss = 0
    for i, (a, b) in enumerate(zip(finals, subs)):
        loss = loss + (a * rec["upstream"][f"w_out{i}"].to(device)).sum() + (b * rec["upstream"][f"w_sub{i}"].to(device)).sum()
    loss.backward()
    torch.cuda.synchronize()
    return net, finals, subs, leaves, time_emb


def check(rec, net, finals, subs, leaves, time_emb, report=None, out_tol=None, grad_tol=None):
    res = {}
    res["time_after"] = _rel(time_emb, rec["time_after"])
    for i, (a, b) in enumerate(zip(finals, rec["finals"])):
        res[f"final{i}"] = _rel(a, b)
    for i, (a, b) in enumerate(zip(subs, rec["subs"])):
        res[f"sub{i}"] = _rel(a, b)
    for k, g in rec["grads_in"].items():
        res["d_" + k] = _rel(leaves[k].grad, g)
    for k, p in net.named_parameters():
        got = p.grad if p.grad is not None else torch.zeros_like(p)
        res["dW_" + k] = _rel(got, rec["grads_param"][k])
    if report is not None:
        report.update(res)
    for k, v in res.items():
        tol = ((out_tol or OUT_TOL) if (k.startswith("final") or k.startswith("sub") or k == "time_after")
               else (grad_tol or GRAD_TOL))
        assert v <= tol, (k, v, res)


@pytest.mark.parametrize("idx", [0, 1, 2, 3])
def test_against_reference_golden(idx):
    rec = torch.load(GOLDEN, weights_only=False)[idx]
    check(rec, *run_product(rec))


def _min_abs_preactivation(sd, args, D, emb, t, it):
    """Smallest |ReLU input| over both stages (fp64 oracle)."""
    m = float("inf")
    for stage, use_coarse in (("c", args.use_coarse_temporal_embedding), ("f", False)):
        if (stage == "c" and args.no_coarse_deform) or (stage == "f" and args.no_fine_deform):
            continue
        n = 30 if use_coarse else (150 if args.no_c2f_temporal_embedding else
                                   deform_oracle.int_lininterp(it, 30, 150, args.c2f_temporal_iter))
        tau = deform_oracle.get_temporal_embed(sd["weight"], t, n)
        x = torch.cat([tau.unsqueeze(0).expand(emb.shape[0], -1), emb], -1)
        for l in range(D):
            x = torch.nn.functional.linear(x if l == 0 else torch.relu(x), sd[f"feature_out_{stage}.{2 * l}.weight"],
                                           sd[f"feature_out_{stage}.{2 * l}.bias"])
            m = min(m, float(x.abs().min()))
        r = torch.relu(x)
        for h in deform_oracle.HEADS:
            z = torch.nn.functional.linear(r, sd[f"{h}_deform_{stage}.1.weight"], sd[f"{h}_deform_{stage}.1.bias"])
            m = min(m, float(z.abs().min()))
    return m


def oracle_record(N, W=128, D=1, it=12345, cam_no=3, t=0.42, seed=3, kink_free=False, device="cpu", **argkw):
    """Record in the golden format but produced by the fp64 oracle (for sizes beyond the fixtures).

    kink_free: hidden biases are set to +-4 / +-40 so every ReLU input is firmly on or off. With millions
    of hidden units a few always sit within rounding distance of the ReLU kink; fp32 (either
    implementation) and fp64 then disagree on relu'(z) for that unit and the corresponding rows of
    the weight gradients differ by O(1/sqrt(N)) -- a property of ReLU, not of the kernels. The small
    golden cases (N=96) test the masking with natural biases; this switch lets the large ragged
    case be compared tightly."""
    from ed3dgs_b200.deformation import deform_network
    d = dict(temporal_embedding_dim=256, gaussian_embedding_dim=32, c2f_temporal_iter=20000, zero_temporal=False,
             no_ds=False, no_dr=False, no_do=False, no_dc=False, no_coarse_deform=False, no_fine_deform=False,
             use_coarse_temporal_embedding=False, no_c2f_temporal_embedding=False, use_anneal=True,
             deform_from_iter=5000)
    d.update(argkw)
    args = NS(**d)
    torch.manual_seed(seed)
    net = deform_network(D=D, W=W, args=args)
    with torch.no_grad():
        net.offsets.copy_(torch.linspace(-0.02, 0.03, 30).view(30, 1))
        net.weight.mul_(30.0)
        if kink_free:
            for name, p in net.named_parameters():
                if name.endswith(".bias") and not name.endswith(".3.bias"):
                    mag = 4.0 if ".0.bias" in name else 40.0   # deeper layers see inputs of magnitude ~4
                    p.copy_(torch.where(torch.arange(p.numel()) % 2 == 0, mag, -mag))
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)  # noqa: E731
    inp = dict(point=r(N, 3), scales=r(N, 3), rotations=r(N, 4), opacity=r(N, 1), sh=r(N, 16, 3),
               embedding=0.5 * r(N, 32))
    up = {}
    for i, shp in enumerate([(N, 3), (N, 3), (N, 4), (N, 1), (N, 16, 3)]):
        up[f"w_out{i}"] = r(*shp)
        up[f"w_sub{i}"] = 0.3 * r(*shp)
    dt = torch.float64
    sd = {k: v.detach().clone().to(device, dt).requires_grad_(True) for k, v in net.state_dict().items()}
    leaves = {k: v.clone().to(device, dt).requires_grad_(True) for k, v in inp.items()}
    time_emb = torch.tensor(t, dtype=dt, device=device).repeat(N, 1)
    out = deform_oracle.forward(sd, args, max(D, 1), 150, leaves["point"], leaves["scales"], leaves["rotations"],
                                leaves["opacity"], time_emb, cam_no, leaves["embedding"], leaves["sh"], it)
    finals, (subs, _o) = out[:5], out[5]
    loss = 0
    for i, (a, b) in enumerate(zip(finals, subs)):
        loss = loss + (a * up[f"w_out{i}"].to(device, dt)).sum() + (b * up[f"w_sub{i}"].to(device, dt)).sum()
    loss.backward()
    if kink_free:
        with torch.no_grad():
            assert _min_abs_preactivation(sd, args, max(D, 1), leaves["embedding"], time_emb[0, 0], it) > 1e-3
    return dict(case=dict(D=D, W=W, it=it, cam_no=cam_no, t=t), args=d,
                state={k: v.detach().clone() for k, v in net.state_dict().items()}, inputs=inp,
                time_after=time_emb.float().cpu(), upstream=up, finals=[x.detach().float().cpu() for x in finals],
                subs=[x.detach().float().cpu() for x in subs],
                grads_in={k: v.grad.float().cpu() for k, v in leaves.items()},
                grads_param={k: (v.grad.float().cpu() if v.grad is not None else torch.zeros_like(v).float().cpu())
                             for k, v in sd.items()})


def test_against_fp64_oracle_ragged_size():
    rec = oracle_record(N=5000 + 37, kink_free=True)
    check(rec, *run_product(rec))


def test_parameter_gradients_at_c3_size_vs_fp64():
    """N = 1 000 003 (C3, ragged by 3): every output, every input gradient and above all every weight / bias gradient
    -- sums over a million Gaussians accumulated across all CTAs with fp32 atomics in arrival order -- against an
    fp64 run of the oracle on the same device. Kink-free biases (see oracle_record), the 1e-4 bound of the small
    cases."""
    rec = oracle_record(N=1_000_003, kink_free=True, seed=5, device="cuda")
    rep = {}
    check(rec, *run_product(rec), report=rep)
    print({k: f"{v:.2e}" for k, v in rep.items() if k.startswith("dW_")})


def test_precision_tf32_mode_stated_tolerance():
    """deform_network.precision = "tf32" (ed3dgs_deform_weights::precision = 1): one tf32 MMA per product, the
    arithmetic torch uses with allow_tf32. Stated tolerance: outputs <= 2e-3, gradients <= 1e-2 relative -- and the
    mode must really be the single-pass one (its error sits far above the 3xTF32 mode's)."""
    rec = oracle_record(N=5000 + 37, kink_free=True)
    rep = {}
    check(rec, *run_product(rec, precision="tf32"), report=rep, out_tol=2e-3, grad_tol=1e-2)
    worst = max(v for k, v in rep.items() if k.startswith("dW_"))
    assert worst > 10 * 2e-5, ("tf32 mode is as accurate as 3xTF32: the switch did not reach the kernels", worst)
    print("tf32 mode:", {k: f"{v:.1e}" for k, v in rep.items() if k.startswith("final") or k.startswith("dW_pos")})


def test_large_natural_biases_outputs_tight_gradients_robust():
    """Natural biases at N=20k: outputs tight; gradients may contain a few ReLU-kink flips, so the
    bound is on the relative L2 error (a flipped unit moves one row of one matrix)."""
    rec = oracle_record(N=20_000 + 3)
    net, finals, subs, leaves, time_emb = run_product(rec)
    for a, b in zip(list(finals) + list(subs), rec["finals"] + rec["subs"]):
        assert _rel(a, b) <= OUT_TOL
    for k, p in net.named_parameters():
        got = p.grad.double().cpu()
        ref = rec["grads_param"][k].double()
        assert float((got - ref).norm() / ref.norm().clamp_min(1e-30)) <= 1e-2, k


def test_heads_disabled_and_single_stage():
    rec = oracle_record(N=700, W=64, D=2, it=100, cam_no=None, kink_free=True, no_ds=True, no_dc=True,
                        no_fine_deform=True)
    check(rec, *run_product(rec))
    rec = oracle_record(N=300, W=128, D=1, it=9000, cam_no=1, kink_free=True, no_coarse_deform=True, no_dr=True)
    check(rec, *run_product(rec))


def test_state_dict_keys_match_reference_names():
    rec = torch.load(GOLDEN, weights_only=False)[0]
    from ed3dgs_b200.deformation import deform_network
    net = deform_network(D=1, W=128, args=NS(**rec["args"]))
    assert set(net.state_dict().keys()) == set(rec["state"].keys())
    assert all(n != "offsets" for n, _ in zip(["x"], net.get_mlp_parameters()))
    assert len(net.get_mlp_parameters()) == len(list(net.parameters())) - 1


def test_full_c3_size_rows_match_fp64_oracle():
    """BASELINE's C3 size (1 M Gaussians, ragged by 3): the network is row-independent, so 4 096 oracle rows are
    scattered into a 1 000 003-row problem (first tile, last ragged tile, random positions) and the product's
    outputs and per-row input gradients at those positions must match the fp64 oracle of the small problem."""
    from ed3dgs_b200.deformation import deform_network
    n_s, N = 4096, 1_000_003
    rec = oracle_record(N=n_s, kink_free=True, seed=11)
    c, args = rec["case"], NS(**rec["args"])
    g = torch.Generator().manual_seed(3)
    pos = torch.cat([torch.arange(0, 300), torch.arange(N - 300, N),
                     torch.randperm(N - 600, generator=g)[:n_s - 600] + 300])
    assert pos.unique().numel() == n_s
    dev = "cuda"
    net = deform_network(D=c["D"], W=c["W"], min_embeddings=30, max_embeddings=150, num_frames=300, args=args)
    net.load_state_dict(rec["state"])
    net = net.to(dev)
    shapes = dict(point=(N, 3), scales=(N, 3), rotations=(N, 4), opacity=(N, 1), sh=(N, 16, 3), embedding=(N, 32))
    leaves = {}
    for k, shp in shapes.items():
        full = torch.randn(*shp, generator=g) * (0.5 if k == "embedding" else 1.0)
        full[pos] = rec["inputs"][k]
        leaves[k] = full.to(dev).requires_grad_(True)
    ups = {}
    for i, shp in enumerate([(N, 3), (N, 3), (N, 4), (N, 1), (N, 16, 3)]):
        for kind, sc in (("w_out", 1.0), ("w_sub", 0.3)):
            full = torch.randn(*shp, generator=g) * sc
            full[pos] = rec["upstream"][f"{kind}{i}"]
            ups[f"{kind}{i}"] = full.to(dev)
    time_emb = torch.tensor(c["t"], device=dev).repeat(N, 1)
    out = net(leaves["point"], leaves["scales"], leaves["rotations"], leaves["opacity"], time_emb, c["cam_no"],
              NS(get_embedding=leaves["embedding"]), None, leaves["sh"], iter=c["it"], num_down_emb_c=30,
              num_down_emb_f=30)
    finals, (subs, _o) = out[:5], out[5]
    loss = 0
    for i, (a, b) in enumerate(zip(finals, subs)):
        loss = loss + (a * ups[f"w_out{i}"]).sum() + (b * ups[f"w_sub{i}"]).sum()
    loss.backward()
    torch.cuda.synchronize()
    p = pos.to(dev)
    for a, b in zip(list(finals) + list(subs), rec["finals"] + rec["subs"]):
        assert _rel(a[p], b) <= OUT_TOL
    for k, gref in rec["grads_in"].items():
        assert _rel(leaves[k].grad[p], gref) <= GRAD_TOL, k
    for k in leaves:
        assert torch.isfinite(leaves[k].grad).all()


@pytest.mark.parametrize("W,D,N", [(256, 1, 3000 + 7), (256, 2, 1500 + 3), (192, 2, 777), (256, 8, 600)])
def test_net_width_up_to_256(W, D, N):
    """deform_network(D=8, W=256) is the reference's constructor default (scene/deformation.py:16): widths above 128 take
    the layered path with the output features of every Linear split into chunks of 128 (two M-side passes) and the
    5*W-long contraction of the trunk's input gradient accumulated head by head."""
    rec = oracle_record(N=N, W=W, D=D, kink_free=True, seed=7)
    check(rec, *run_product(rec))
