"""CPU: the deformation oracle (oracle/deform_oracle.py) against the golden vectors produced by the
reference's own deform_network, and -- when the reference tree is present -- against that module
run live. Tolerance: fp32 CPU matmul reassociation only (1e-5 relative to the tensor's max)."""
import os
import sys
from types import SimpleNamespace as NS

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import deform_oracle  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden", "deform_golden.pt")


def _rel(a, b):
    scale = float(b.abs().max())
    return float((a - b).abs().max()) / (scale if scale > 0 else 1.0)


def run_oracle(rec, dtype=torch.float32):
    c = rec["case"]
    args = NS(**rec["args"])
    sd = {k: v.clone().to(dtype).requires_grad_(True) for k, v in rec["state"].items()}
    leaves = {k: v.clone().to(dtype).requires_grad_(True) for k, v in rec["inputs"].items()}
    N = leaves["point"].shape[0]
    time_emb = torch.tensor(c["t"], dtype=dtype).repeat(N, 1)
    out = deform_oracle.forward(sd, args, max(c["D"], 1), 150, leaves["point"], leaves["scales"], leaves["rotations"],
                                leaves["opacity"], time_emb, c["cam_no"], leaves["embedding"], leaves["sh"], c["it"])
    finals, (subs, _o) = out[:5], out[5]
    loss = 0
    for i, (a, b) in enumerate(zip(finals, subs)):
        loss = loss + (a * rec["upstream"][f"w_out{i}"].to(dtype)).sum() + (b * rec["upstream"][f"w_sub{i}"].to(dtype)).sum()
    loss.backward()
    return finals, subs, leaves, sd, time_emb


@pytest.mark.parametrize("idx", [0, 1, 2, 3])
def test_oracle_matches_reference_golden(idx):
    rec = torch.load(GOLDEN, weights_only=False)[idx]
    finals, subs, leaves, sd, time_emb = run_oracle(rec)
    assert torch.allclose(time_emb, rec["time_after"])
    for a, b in zip(finals, rec["finals"]):
        assert _rel(a.detach(), b) < 1e-5
    for a, b in zip(subs, rec["subs"]):
        assert _rel(a.detach(), b) < 1e-5
    for k, g in rec["grads_in"].items():
        assert _rel(leaves[k].grad, g) < 1e-5, k
    for k, g in rec["grads_param"].items():
        got = sd[k].grad if sd[k].grad is not None else torch.zeros_like(g)
        assert _rel(got, g) < 2e-5, k


# shapes beyond the committed fixtures: the reference constructor's own default (D=8, W=256, scene/deformation.py:16) and a
# wide two-layer trunk -- the shapes tests/test_deform_gpu.py::test_net_width_up_to_256 checks the kernels against this
# oracle at, so the oracle is pinned there on the reference module itself
LIVE_CASES = [None,
              dict(name="ctor_default_d8_w256", args={}, D=8, W=256, it=12345, cam_no=3, t=0.42),
              dict(name="w256_d2_nodc", args=dict(no_dc=True), D=2, W=256, it=600, cam_no=None, t=0.7)]


@pytest.mark.skipif(not os.path.exists("/root/reference/scene/deformation.py"), reason="reference tree absent")
@pytest.mark.parametrize("case", LIVE_CASES, ids=lambda c: "fixture_case_2" if c is None else c["name"])
def test_oracle_matches_live_reference(case):
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_deform_golden as mk
    mod = mk.load_reference()
    rec = mk.run_case(mod, mk.CASES[2] if case is None else case, N=40)
    finals, subs, leaves, sd, _ = run_oracle(rec)
    for a, b in zip(finals, rec["finals"]):
        assert _rel(a.detach(), b) < 1e-5
    for k, g in rec["grads_in"].items():
        assert _rel(leaves[k].grad, g) < 1e-5, k
