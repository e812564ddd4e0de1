"""bench.py's work models against the figures SURVEY.md 8(d) / BASELINE.md 5 state, and the CPU reference leg
(oracle port) end to end at a tiny size."""
import os
import sys

import pytest
import torch

import conftest  # noqa: F401
sys.path.insert(0, conftest.ROOT)
import bench  # noqa: E402
from ed3dgs_b200.renderer import default_hyper  # noqa: E402


def test_mlp_flop_model_is_the_minimal_formulation():
    cfg = dict(P=1_000_000, W=1604, H=1100)
    work = bench.algorithmic_work(cfg, default_hyper(), R=8_000_000, visible_frac=1.0)
    per_pass = work["fused_fwd_kernel"][1] / cfg["P"]
    assert per_pass == 2 * 187_136                       # two stages
    assert abs(3 * per_pass - 1.123e6) < 1e3            # forward + input gradients + weight gradients
    assert work["tc_wgrad_kernel"] == work["fused_bwd_kernel"] == work["fused_fwd_kernel"]
    assert work["fused_fwd_kernel"][0] == "tensor"
    # rasterizer: the per-kernel figures add up to B_ras = P(268 + 770 V) + 240 R + 124 px
    R, px, P = 8_000_000, cfg["W"] * cfg["H"], cfg["P"]
    total = sum(work[k][1] for k in bench.RASTER_GROUPS)
    assert abs(total - (P * (268 + 770) + 240 * R + 124 * px)) / total < 1e-9


def test_technicolor_workload_shape():
    from ed3dgs_b200 import synthetic
    h = bench.workload_hyper(synthetic.CONFIGS["C5"])
    assert (h.defor_depth, h.no_dc, h.min_embeddings, h.max_embeddings) == (0, True, 5, 25)
    m = bench.mlp_shape(h)
    assert (m["nh"], m["k_on"], m["D"]) == (4, 11, 1)


def test_cpu_reference_leg_runs_full_pipeline_small():
    cfg = dict(P=3000, W=96, H=64, sh_degree=3, frames=150, focal_full=8200.0, full_w=3208 * 96 / 1604)
    ref = bench.CpuReference(cfg, 1.0)
    med, ts = ref.run(steps=2, warmup=1, budget_s=60.0)
    assert len(ts) == 2 and med[0] > 0 and ref.R > 0
    assert all(v.grad is not None and torch.isfinite(v.grad).all() for k, v in ref.sd.items()
               if k.startswith("pos_deform_f"))
    assert "no extrapolation" in ref.describe(med, ts, 2)


def test_both_arms_print_the_same_config_object():
    """The driver compares the two arms' `config` dicts: nothing a run observes (R, visible fraction, host details)
    may sit in it."""
    from ed3dgs_b200 import synthetic
    c = bench.workload_config("C3", synthetic.CONFIGS["C3"], 0.3)
    assert set(c) == {"workload", "splat_scale", "l2_policy"} and "1000000 Gaussians" in c["workload"]
    assert c == bench.workload_config("C3", synthetic.CONFIGS["C3"], 0.3)
    d = bench.workload_detail(synthetic.CONFIGS["C3"], 8_800_000, 1, 0.65, host="x")
    assert d["num_rendered"] == 8_800_000 and d["visible_fraction"] == 0.65 and d["host"] == "x"
