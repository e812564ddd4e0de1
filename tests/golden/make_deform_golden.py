"""Generates tests/golden/deform_golden.pt by running the REFERENCE's own deform_network
(/root/reference/scene/deformation.py, imported by file path with tkinter stubbed and .cuda()
neutralised) on small seeded inputs, forward + backward, on CPU. Run in the dev container only."""
import importlib.util
import os
import sys
import types
from types import SimpleNamespace as NS

import torch

REF = "/root/reference/scene/deformation.py"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "deform_golden.pt")


def load_reference():
    sys.modules.setdefault("tkinter", types.SimpleNamespace(W=None))
    spec = importlib.util.spec_from_file_location("ref_deformation", REF)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def make_args(**kw):
    d = dict(temporal_embedding_dim=256, gaussian_embedding_dim=32, c2f_temporal_iter=20000, zero_temporal=False,
             no_ds=False, no_dr=False, no_do=False, no_dc=False, no_coarse_deform=False, no_fine_deform=False,
             use_coarse_temporal_embedding=False, no_c2f_temporal_embedding=False, use_anneal=True,
             deform_from_iter=5000, net_width=128, defor_depth=1)
    d.update(kw)
    return NS(**d)


CASES = [
    dict(name="nersemble_iter0", args={}, D=1, W=128, it=0, cam_no=2, t=0.37),
    dict(name="nersemble_iter20000", args={}, D=1, W=64, it=20000, cam_no=None, t=0.81),
    dict(name="depth2_w64_nodc", args=dict(no_dc=True, no_do=True, use_coarse_temporal_embedding=True), D=2, W=32,
         it=7000, cam_no=0, t=0.05),
    dict(name="technicolor_like", args=dict(no_dc=True, use_anneal=False), D=0, W=64, it=300, cam_no=1, t=0.999),
]


def run_case(mod, c, N=96):
    torch.manual_seed(1234)
    args = make_args(**c["args"])
    net = mod.deform_network(D=c["D"], W=c["W"], min_embeddings=30, max_embeddings=150, num_frames=300, args=args)
    with torch.no_grad():
        net.offsets.copy_(torch.linspace(-0.02, 0.03, 30).view(30, 1))
        net.weight.mul_(30.0)
    g = torch.Generator().manual_seed(77)
    r = lambda *s: torch.randn(*s, generator=g)  # noqa: E731
    inp = dict(point=r(N, 3), scales=r(N, 3), rotations=r(N, 4), opacity=r(N, 1), sh=r(N, 16, 3),
               embedding=0.5 * r(N, 32))
    leaves = {k: v.clone().requires_grad_(True) for k, v in inp.items()}
    pc = NS(get_embedding=leaves["embedding"])
    time_emb = torch.tensor(c["t"]).repeat(N, 1)
    torch.Tensor.cuda = lambda self, *a, **k: self  # the reference hard-codes .cuda() at :61
    out = net(leaves["point"], leaves["scales"], leaves["rotations"], leaves["opacity"], time_emb, c["cam_no"], pc,
              None, leaves["sh"], iter=c["it"], num_down_emb_c=30, num_down_emb_f=30)
    finals, (subs, _orig) = out[:5], out[5]
    gw = torch.Generator().manual_seed(5)
    loss = 0
    up = {}
    for i, (a, b) in enumerate(zip(finals, subs)):
        wa, wb = torch.randn(a.shape, generator=gw), 0.3 * torch.randn(b.shape, generator=gw)
        up[f"w_out{i}"], up[f"w_sub{i}"] = wa, wb
        loss = loss + (a * wa).sum() + (b * wb).sum()
    loss.backward()
    rec = dict(name=c["name"], case=c, args=vars(args), state={k: v.detach().clone() for k, v in net.state_dict().items()},
               inputs=inp, time_after=time_emb.detach().clone(), upstream=up,
               finals=[x.detach().clone() for x in finals], subs=[x.detach().clone() for x in subs],
               grads_in={k: v.grad.detach().clone() for k, v in leaves.items()},
               grads_param={k: (p.grad.detach().clone() if p.grad is not None else torch.zeros_like(p))
                            for k, p in net.named_parameters()})
    return rec


def main():
    mod = load_reference()
    recs = [run_case(mod, c) for c in CASES]
    torch.save(recs, OUT)
    print("wrote", OUT, os.path.getsize(OUT) // 1024, "KiB")


if __name__ == "__main__":
    main()
