#!/usr/bin/env python
"""bench.py -- train fwd+bwd step of the E-D3DGS hot path (deformation MLP + RaDe-GS rasterizer).

  python bench.py --gpus N --steps K --warmup W            (our CUDA path; torchrun for N > 1)
  python bench.py --impl reference --gpus N --steps K --warmup W   (CPU oracle port, rank 0 only)

A step = one view per rank: render() (deformation network -> activations -> rasterizer) ->
loss -> backward (rasterizer backward -> deformation backward) -> gradient all-reduce (N > 1).
Workload: BASELINE.json configs[2] = "C3" (1M Gaussians, 1604x1100, depth+normal outputs).
Prints ONE JSON line on rank 0. See DESIGN.md section "Measurement" for every field.
"""
import argparse
import json
import math
import os
import subprocess
import sys
import threading
import time
from types import SimpleNamespace as NS

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "e-d3dgs_b200"))

import torch  # noqa: E402

METRIC = "train fwd+bwd views/s (1M Gaussians, 1604x1100, depth+normal outputs); ms_per_step = ms/iter"
UNIT = "views/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="C3")
    ap.add_argument("--splat-scale", type=float, default=0.3)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-gpu-reference", action="store_true")
    return ap.parse_args()


def load_peaks():
    """HBM GB/s and dense bf16 TFLOP/s: MEASURED_PEAKS.json (driver-written) else the profiling recipe's fallback.
    The tf32 tensor peak the MLP rooflines use is measured in this run (measure_tf32_peak) and falls back to half
    the sustained bf16 figure."""
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm_gbs=d["hbm_gbs"], bf16_tflops=d.get("bf16_tflops_sustained", d["bf16_tflops"]), source="measured")
    return dict(hbm_gbs=6650.0, bf16_tflops=1400.0, source="fallback")


def measure_tf32_peak(device, n=8192, iters=10):
    """cuBLAS tf32 throughput on this GPU, the denominator of the MLP tensor rooflines: torch.matmul of fp32
    n^3 with TF32 allowed, best of `iters` (burst) -- measured outside every timed region; the flag is restored."""
    keep = torch.backends.cuda.matmul.allow_tf32
    try:
        torch.backends.cuda.matmul.allow_tf32 = True
        a = torch.randn(n, n, device=device)
        b = torch.randn(n, n, device=device)
        for _ in range(3):
            a @ b
        best = 1e30
        for _ in range(iters):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            a @ b
            e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        return 2.0 * n ** 3 / (best * 1e-3) / 1e12
    except Exception:  # noqa: BLE001
        return None
    finally:
        torch.backends.cuda.matmul.allow_tf32 = keep
        torch.cuda.empty_cache()


class SmiClockSampler:
    """Same report as ClockSampler, sampled by an `nvidia-smi -lms 250` child process (the profiling recipe's own
    line) started before the warm-up, so that its driver attach is outside the timed region; its csv rows carry
    wall-clock timestamps and only rows between mark_begin() and mark_end() are reported."""

    def __init__(self, gpu_index):
        import subprocess
        import tempfile
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = int(vis.split(",")[gpu_index]) if vis and vis.split(",")[0].isdigit() else gpu_index
        self.out = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        self.t0 = self.t1 = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(idx), "--query-gpu=timestamp,clocks.sm,clocks.max.sm,"
                 "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
                 "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap",
                 "--format=csv,noheader,nounits", "-lms", "250"], stdout=self.out, stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def mark_begin(self):
        self.t0 = time.time()

    def mark_end(self):
        self.t1 = time.time()

    def stop(self):
        import datetime
        if self.proc is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:  # noqa: BLE001
            self.proc.kill()
        self.out.flush()
        self.out.seek(0)
        rows = []
        for ln in self.out.read().splitlines():
            f = [x.strip() for x in ln.split(",")]
            try:
                ts = datetime.datetime.strptime(f[0], "%Y/%m/%d %H:%M:%S.%f").timestamp()
                rows.append((ts, float(f[1]), float(f[2]), f[3:7]))
            except (ValueError, IndexError):
                continue
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        inside = [r for r in rows if self.t0 is not None and self.t0 <= r[0] <= (self.t1 or 1e30)]
        if not inside and rows:
            inside = sorted(rows, key=lambda r: abs(r[0] - (self.t0 or 0)))[:1]
        if not inside:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["no nvidia-smi samples"])
        sm = sorted(r[1] for r in inside)
        reasons = sorted({n for r in inside for n, v in zip(names, r[3]) if v.lower().startswith("active")})
        return dict(sm_mhz=sm[len(sm) // 2], sm_max_mhz=inside[0][2], reasons=reasons, samples=len(sm),
                    how="nvidia-smi -lms 250 child")


class ClockSampler:
    """SM clocks / throttle reasons sampled every 250 ms through NVML (pynvml) from a thread of this process
    (every NVML query takes a driver lock that delays kernel launches by a millisecond or more, so the polling
    interval is part of what the timed region measures: 100 ms cost 0.8 ms per 16 ms step).
    NVML is initialised before the warm-up so no driver attach happens inside the timed region (an
    `nvidia-smi -lms` child started there stalled launches for hundreds of ms on some boxes); only samples
    taken between mark_begin() and mark_end() are reported."""
    REASONS = ((0x8, "hw_slowdown"), (0x40, "hw_thermal_slowdown"), (0x20, "sw_thermal_slowdown"), (0x4, "sw_power_cap"))

    def __init__(self, gpu_index):
        self.samples = []
        self.t0 = self.t1 = None
        self.stop_flag = threading.Event()
        self.h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            vis = os.environ.get("CUDA_VISIBLE_DEVICES")
            idx = int(vis.split(",")[gpu_index]) if vis and vis.split(",")[0].isdigit() else gpu_index
            self.h = pynvml.nvmlDeviceGetHandleByIndex(idx)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
            self.thread = threading.Thread(target=self._poll, daemon=True)
            self.thread.start()
        except Exception:  # noqa: BLE001
            self.h = None

    def _poll(self):
        while not self.stop_flag.wait(0.25):
            try:
                mhz = float(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                self.samples.append((time.perf_counter(), mhz, mask))
            except Exception:  # noqa: BLE001
                pass

    def mark_begin(self):
        self.t0 = time.perf_counter()

    def mark_end(self):
        self.t1 = time.perf_counter()

    def stop(self):
        if self.h is None:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvml unavailable"])
        self.stop_flag.set()
        inside = [s for s in self.samples if self.t0 is not None and self.t0 <= s[0] <= (self.t1 or 1e30)]
        if not inside:  # region shorter than one polling interval: nearest sample
            inside = sorted(self.samples, key=lambda s: abs(s[0] - (self.t0 or 0)))[:1]
        sm = sorted(s[1] for s in inside)
        reasons = sorted({name for s in inside for bit, name in self.REASONS if s[2] & bit})
        return dict(sm_mhz=(sm[len(sm) // 2] if sm else None), sm_max_mhz=self.max_mhz, reasons=reasons,
                    samples=len(sm))


def workload_hyper(cfg):
    """ModelHiddenParams of the dataset config a workload stands for: nersemble / dynerf
    defaults for C1-C4, technicolor (arguments/technicolor/default.py:6-24: defor_depth 0, no SH head, 5/25
    temporal embeddings) for C5."""
    from ed3dgs_b200.renderer import default_hyper
    kw = dict(total_num_frames=cfg["frames"])
    if cfg.get("family") == "technicolor":
        kw.update(defor_depth=0, no_dc=True, min_embeddings=5, max_embeddings=25)
    return default_hyper(**kw)


def build_scene(cfg, splat_scale, device, n_cams=16):
    from ed3dgs_b200 import synthetic
    from ed3dgs_b200.renderer import GaussianScene
    g = synthetic.make_gaussians(cfg["P"], splat_scale=splat_scale, device="cpu")
    hyper = workload_hyper(cfg)
    torch.manual_seed(synthetic.SEED)
    scene = GaussianScene(g, hyper, sh_degree=cfg["sh_degree"]).to(device)
    with torch.no_grad():
        scene._deformation.weight.mul_(30.0)  # a trained-looking temporal table instead of ~1e-3 noise
    focal = cfg["focal_full"] * cfg["W"] / cfg["full_w"]
    cams = [synthetic.make_camera(k, n_cams, cfg["W"], cfg["H"], focal, frame=(7 * k) % cfg["frames"],
                                  frames=cfg["frames"], device="cpu") for k in range(n_cams)]
    return scene, cams


def cam_to(cam, device, non_blocking=True):
    c = NS(**vars(cam))
    for k in ("world_view_transform", "full_proj_transform", "camera_center"):
        setattr(c, k, getattr(cam, k).to(device, non_blocking=non_blocking))
    return c


def profile_report():
    import ctypes
    from ed3dgs_b200 import _lib
    lib = _lib.lib()
    lib.ed3dgs_profile_report.restype = ctypes.c_int
    lib.ed3dgs_profile_report.argtypes = [ctypes.c_char_p, ctypes.c_size_t]
    buf = ctypes.create_string_buffer(1 << 16)
    lib.ed3dgs_profile_report(buf, len(buf))
    out = {}
    for ln in buf.value.decode().splitlines():
        name, ms, n = ln.rsplit(" ", 2)
        out[name] = (float(ms), int(n))
    return out


def profile_pairs():
    import ctypes
    from ed3dgs_b200 import _lib
    lib = _lib.lib()
    out = (ctypes.c_ulonglong * 2)()
    lib.ed3dgs_profile_pairs.restype = ctypes.c_int
    lib.ed3dgs_profile_pairs.argtypes = [ctypes.POINTER(ctypes.c_ulonglong)]
    return (int(out[0]), int(out[1])) if lib.ed3dgs_profile_pairs(out) == 0 else None


def profile_enable(on):
    from ed3dgs_b200 import _lib
    _lib.lib().ed3dgs_profile_enable(1 if on else 0)


# (the weight gradients run as tc_wgrad_ts_kernel -- M-side operand in tensor memory -- in the product build; the
# shared-memory tc_wgrad_kernel only behind ED3DGS_WG_TS=0: the library's scope carries whichever name ran)
MLP_GROUPS = ("fused_fwd_kernel", "fused_bwd_kernel", "tc_wgrad_ts_kernel", "tc_wgrad_kernel",
              "tc_rows_kernel<fwd>", "tc_rows_kernel<dgrad>")
RASTER_GROUPS = ("preprocess_fwd", "scan_tiles", "emit_keys", "radix_sort_pairs", "tile_ranges", "render_fwd",
                 "blend_bwd", "gaussian_bwd")


def mlp_shape(hyper):
    heads = [True, not hyper.no_ds, not hyper.no_dr, not hyper.no_do, not hyper.no_dc]
    k_on = sum(k for k, on in zip((3, 3, 4, 1, 48), heads) if on)
    stages = (0 if hyper.no_coarse_deform else 1) + (0 if hyper.no_fine_deform else 1)
    return dict(W=hyper.net_width, nh=sum(heads), k_on=k_on, stages=stages, ge=hyper.gaussian_embedding_dim,
                D=max(hyper.defor_depth, 1))


def algorithmic_work(cfg, hyper, R, visible_frac):
    """Per-step algorithmic work of each kernel group, exactly as SURVEY.md 8(d) / BASELINE.md 5 define it.
    Deformation network: TENSOR model on the minimal flops -- per stage and pass (forward, input gradients, weight
    gradients) 2*(32*W + (D-1)*W^2 + H_on*W^2 + W*k_on) flop per Gaussian (187 136 at the defaults; 1.123 MFLOP
    per Gaussian for forward + backward of both stages); the kernels execute 3 tf32 MMAs per algorithmic product
    (3xTF32) and are reported against this 1x figure. Rasterizer kernels: HBM bytes."""
    P, px = cfg["P"], cfg["W"] * cfg["H"]
    m = mlp_shape(hyper)
    W, nh, k_on, stages, ge, D = m["W"], m["nh"], m["k_on"], m["stages"], m["ge"], m["D"]
    flop_pass = 2 * (ge * W + (D - 1) * W * W + nh * W * W + W * k_on) * stages * P
    V = visible_frac
    work = {name: ("tensor", flop_pass) for name in MLP_GROUPS}
    work.update({
        "preprocess_fwd": ("hbm", P * (240 + 143 * V)),
        "scan_tiles": ("hbm", 8 * P),
        "emit_keys": ("hbm", 20 * P + 12 * R),
        "radix_sort_pairs": ("hbm", 24 * R),          # one read + one write of the 12-byte pairs
        "tile_ranges": ("hbm", 8 * R),
        "render_fwd": ("hbm", 68 * R + 56 * px),
        "blend_bwd": ("hbm", 128 * R + 68 * px),
        "gaussian_bwd": ("hbm", (367 + 260) * V * P),
        # fused activations: forward reads + writes 32 B per Gaussian, backward reads 64 B and writes 32 B
        "activations": ("hbm", (64 + 96) * P),
        # fused image losses: image + target read, three derivative maps written / read back, gradient written
        "ssim_l1_forward": ("hbm", 3 * px * 4 * (2 + 3)),
        "ssim_l1_backward": ("hbm", 3 * px * 4 * (2 + 3 + 1)),
        "depth_normal_forward": ("hbm", px * 4 * 5),
        "depth_normal_backward": ("hbm", px * 4 * (5 + 5)),
    })
    return work


def mlp_hbm_model(hyper, P):
    """HBM side of the deformation kernels, bytes per step. `algorithmic` is SURVEY 8(d)'s figure for a fully fused
    network (forward: 364 B in + 236 B out + 236 B extras per Gaussian for both stages together; backward: the
    upstream gradients of the outputs and of the coarse-stage extras in, the embedding gradient out; the weight
    gradients need no per-Gaussian bytes of their own). `design` is what THIS design moves: the forward leaves the
    post-ReLU activations feature-major per 128-Gaussian tile (zT, hT) and the ReLU masks, the backward chain leaves
    dzT / dhT / GT, and the weight-gradient kernel contracts those matrices over the Gaussians."""
    m = mlp_shape(hyper)
    W, nh, k_on, stages, ge = m["W"], m["nh"], m["k_on"], m["stages"], m["ge"]
    out_b = 4 * k_on
    masks = (nh + 1) * (W // 8)
    fwd = stages * (4 * ge + out_b + out_b + 4 * W * (nh + 1) + masks)
    bwd = stages * (2 * out_b + masks + 4 * W * (nh + 1) + 4 * 64 + 4 * ge) + 4 * ge
    wg = stages * (2 * 4 * W * nh + 2 * 4 * W + 4 * 64 + 4 * ge)
    return {
        "fused_fwd_kernel": {"algorithmic": (4 * ge + 3 * out_b + 4 * ge) * P if stages else 0, "design": fwd * P},
        "fused_bwd_kernel": {"algorithmic": (2 * out_b + 4 * ge) * P, "design": bwd * P},
        "tc_wgrad_ts_kernel": {"algorithmic": 0, "design": wg * P},
        "tc_wgrad_kernel": {"algorithmic": 0, "design": wg * P},
    }


TRAFFIC_FILE = "profiles/r03_traffic.json"   # written by tests/make_profiles.py from gpurun_out/r03 (tests/gpu_round.sh)
SORT_DIGIT_BITS = 8
SORT_LAUNCHES_PER_PASS, SORT_LAUNCHES_FIXED = 1, 2   # csrc/radix_sort.cu: one-sweep kernel per pass + histogram + digit scan


def launches_per_step(prof, steps, sort_passes):
    """Kernel launches of THIS library per step, from its own event scopes (a scope holding
    several launches is weighted; the dozen tiny unscoped kernels are added as a constant)."""
    mult = {"scan_tiles": 3, "radix_sort_pairs": SORT_LAUNCHES_PER_PASS * sort_passes + SORT_LAUNCHES_FIXED}
    n = 0
    for name, (_ms, cnt) in prof.items():
        n += cnt * mult.get(name, 1)
    return int(round(n / max(steps, 1))) + 16


# ------------------------------------------------------------------------------------------------
# CPU legs: the oracle port (oracle/deform_oracle.py + oracle/oracle_raster.cpp) on the host cores.
# ONE helper serves both the GPU arm's `cpu_baseline` and `--impl reference`, so the two figures are the same
# measurement: the FULL workload (same seeded scene, same camera, same loss as the GPU arm), no extrapolation.
class CpuReference:
    """The reference path restated for the host cores: torch-CPU deformation oracle (fp32, all threads) ->
    activations -> C++/OpenMP rasterizer oracle -> the reference's torch losses -> backward through all of it."""

    def __init__(self, cfg, splat_scale):
        sys.path.insert(0, os.path.join(ROOT, "oracle"))
        import deform_oracle
        import loss_oracle
        import raster_oracle
        from ed3dgs_b200 import synthetic
        torch.set_num_threads(os.cpu_count() or 1)
        self.mods = (deform_oracle, loss_oracle, raster_oracle)
        self.cfg = cfg
        scene, cams = build_scene(cfg, splat_scale, "cpu")
        self.hyper = scene.hyper
        self.sd = {k: v.detach().clone().requires_grad_(True) for k, v in scene._deformation.state_dict().items()}
        self.g = dict(xyz=scene._xyz.detach(), scaling=scene._scaling.detach(), rotation=scene._rotation.detach(),
                      opacity=scene._opacity.detach(), embedding=scene._embedding.detach(),
                      features=torch.cat((scene._features_dc, scene._features_rest), 1).detach())
        self.tongue = scene.tongue_class
        self.cams = cams
        self.sh_degree = cfg["sh_degree"]
        self.gt = torch.rand(3, cfg["H"], cfg["W"], generator=torch.Generator().manual_seed(11))
        self.R = 0

    def step(self, i=0):
        """One fwd+bwd step at the full size. Returns (total, network, rasterizer+loss) seconds."""
        deform_oracle, loss_oracle, raster_oracle = self.mods
        from ed3dgs_b200 import synthetic
        hyper, cfg = self.hyper, self.cfg
        cam = self.cams[i % len(self.cams)]
        P = cfg["P"]
        rs = synthetic.raster_settings(cam, self.sh_degree, kernel_size=0.0, require_coord=False, require_depth=True)
        t0 = time.perf_counter()
        for v in self.sd.values():
            v.grad = None
        leaves = {k: v.clone().requires_grad_(True) for k, v in self.g.items()}
        time_emb = torch.tensor(cam.time).repeat(P, 1)
        out = deform_oracle.forward(self.sd, hyper, max(hyper.defor_depth, 1), hyper.max_embeddings, leaves["xyz"],
                                    leaves["scaling"], leaves["rotation"], leaves["opacity"], time_emb, cam.uid % 30,
                                    leaves["embedding"], leaves["features"], 20000,
                                    num_down_emb_c=hyper.min_embeddings, num_down_emb_f=hyper.min_embeddings)
        means, scales = out[0], torch.exp(out[1])
        rots, opac = torch.nn.functional.normalize(out[2]), torch.sigmoid(out[3])
        t1 = time.perf_counter()
        orc = raster_oracle.OracleRasterizer()
        self.R, o = orc.forward(rs, means, opac, self.tongue, shs=out[4], scales=scales, rotations=rots)
        names = ("color", "depth", "mdepth", "alpha", "normal")
        leaf = {k: o[k].clone().requires_grad_(True) for k in names}
        loss = (loss_oracle.l1_loss(leaf["color"], self.gt) + 0.2 * (1 - loss_oracle.ssim(leaf["color"], self.gt)) / 2
                + 0.05 * loss_oracle.depth_normal_loss(cam, leaf["depth"], leaf["mdepth"], leaf["normal"])
                + 0.01 * leaf["alpha"].mean())
        gr = dict(zip(names, torch.autograd.grad(loss, [leaf[k] for k in names])))
        gb = orc.backward(gr)
        t2 = time.perf_counter()
        torch.autograd.backward(
            [means, scales, rots, opac, out[4]],
            [gb["dL_dmeans3D"], gb["dL_dscales"], gb["dL_drotations"], gb["dL_dopacity"], gb["dL_dsh"]])
        t3 = time.perf_counter()
        return t3 - t0, (t1 - t0) + (t3 - t2), t2 - t1

    def run(self, steps, warmup, budget_s):
        """`warmup` untimed full steps (at least one, so thread pools, allocator and page cache are warm), then up
        to `steps` timed ones; stops early once `budget_s` seconds of timed work are spent (never below 1 step)."""
        for i in range(max(1, warmup)):
            self.step(i)
        ts, spent = [], 0.0
        for i in range(max(1, steps)):
            t = self.step(i)
            ts.append(t)
            spent += t[0]
            if spent >= budget_s:
                break
        med = sorted(ts)[len(ts) // 2]
        return med, ts

    def describe(self, med, ts, asked_steps):
        c = self.cfg
        return (f"FULL workload, no extrapolation: {c['P']} Gaussians at {c['W']}x{c['H']}, R={self.R}, same scene / "
                f"cameras / loss as the GPU arm; torch-CPU fp32 deformation oracle + C++/OpenMP rasterizer oracle on "
                f"{os.cpu_count()} threads; {len(ts)} timed steps of {asked_steps} asked (time-capped) after a full "
                f"warm-up step, median {med[0]:.2f} s (network {med[1]:.2f} s, rasterizer + losses {med[2]:.2f} s)")


def run_reference_arm(args, rank, world):
    """--impl reference: the CPU port of the reference path on rank 0: full-size steps, honouring --steps/--warmup
    up to a wall-clock cap (BENCH_REF_BUDGET_S, default 120 s of timed steps; warm-up capped at 1 full step)."""
    if rank != 0:
        return
    from ed3dgs_b200 import synthetic
    cfg = synthetic.CONFIGS[args.workload]
    budget = float(os.environ.get("BENCH_REF_BUDGET_S", "120"))
    ref = CpuReference(cfg, args.splat_scale)
    med, ts = ref.run(args.steps, min(max(args.warmup, 1), 1), budget)
    mean = sum(t[0] for t in ts) / len(ts)
    sample = ref.describe(med, ts, args.steps)
    line = {"impl": "reference", "metric": METRIC, "value": 1.0 / mean, "unit": UNIT, "n_gpus": args.gpus,
            "steps": len(ts), "warmup": 1, "ms_per_step": mean * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args.workload, cfg, args.splat_scale),
            "workload_detail": workload_detail(cfg, ref.R, 1, view=(len(ts) - 1) % len(ref.cams)),
            "cpu_baseline": {"value": 1.0 / mean, "unit": UNIT, "cores": os.cpu_count(), "kind": "port", "sample": sample},
            "e2e": {"value": 1.0 / mean, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def workload_config(name, cfg, splat_scale):
    """The `config` object BOTH arms print, key for key and value for value (the driver compares the two dicts):
    only what defines the workload. What a run observed about it (R of the view it rendered last, visible fraction,
    host and multi-GPU details) goes into the sibling key `workload_detail`."""
    return {"workload": f"{name}: {cfg['P']} Gaussians, {cfg['W']}x{cfg['H']}, SH degree {cfg['sh_degree']}, "
                        f"variant <coord=0,depth=1,normal=1>, kernel_size 0, one view per rank per step",
            "splat_scale": splat_scale,
            "l2_policy": "inputs larger than L2: per-step working set (Gaussian records, keys / lists, MLP "
                         "activations) exceeds the 126 MB L2 many times over"}


def workload_detail(cfg, R, world, vis=None, **extra):
    d = {"num_rendered": R, "R_over_P": R / cfg["P"], "views_per_step": world}
    if vis is not None:
        d["visible_fraction"] = vis
    d.update(extra)
    return d


# ------------------------------------------------------------------------------------------------
def run_ours(args, rank, local, world):
    import torch.distributed as dist
    from ed3dgs_b200 import _lib, dp, synthetic
    from ed3dgs_b200.renderer import render, training_loss
    _lib.lib()  # fail loudly if the CUDA library is missing
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: the product path has no CPU fallback")
    device = torch.device(f"cuda:{local}")
    torch.cuda.set_device(device)
    cfg = synthetic.CONFIGS[args.workload]
    scene, cams = build_scene(cfg, args.splat_scale, device)
    hyper = scene.hyper
    # gradients: set to None before every step like the reference's zero_grad(set_to_none=True) (train.py:448); for N > 1
    # all of them are reduced in one NCCL group (ed3dgs_b200/dp.py: GradSync)
    # N > 1: the 59 of 91 per-Gaussian gradient floats that are final before the network's backward starts are exchanged
    # next to it (dp.OverlapSync). Default transport "ce": peer-to-peer copy-engine transfers between symmetric-memory
    # buffers (8 GPUs, C3: 631 views/s against 591 with one NCCL exchange at the end of the step;
    # NCCL next to the backward is slower than either, it needs far more CTAs than the persistent MLP kernels can spare).
    # BENCH_OVERLAP = ce | nccl | 0. If symmetric memory cannot be set up the bench falls back to the single exchange.
    ov_mode = os.environ.get("BENCH_OVERLAP", "ce")
    ov_mode = "nccl" if ov_mode == "1" else ov_mode
    overlap = world > 1 and ov_mode != "0"
    transport = "ce" if ov_mode == "ce" else "nccl"
    sm_reserve = int(os.environ.get("BENCH_SM_RESERVE", "8" if transport == "nccl" else "2"))
    arena = dp.OverlapSync(list(scene.parameters()),
                           early=[scene._xyz, scene._features_dc, scene._features_rest, scene._scaling, scene._rotation,
                                  scene._opacity], sm_reserve=sm_reserve, transport=transport,
                           embedding=scene._embedding if os.environ.get("BENCH_CE_EMB", "1") != "0" else None,
                           sm_reserve_last=int(os.environ.get("BENCH_SM_RESERVE_LAST", "8")))
    overlap_note = None
    if overlap:
        try:
            overlap = arena.enable()
        except Exception as ex:  # noqa: BLE001   (no symmetric memory on this box: one NCCL exchange at the end of the step)
            overlap_note = f"{transport} exchange unavailable ({type(ex).__name__}: {ex}); single NCCL exchange"
            arena.transport, arena.ce, arena.active = "nccl", None, False
            overlap = False
        if world > 1:   # every rank must have made the same choice
            flag = torch.tensor([1 if overlap else 0], device=device)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if overlap and int(flag) == 0:
                arena.disable()
                overlap, overlap_note = False, "a rank could not set the overlapped exchange up; single NCCL exchange"
    pipe = NS(debug=False)
    bg = torch.zeros(3, device=device)
    H, W = cfg["H"], cfg["W"]
    # a few distinct target images in pinned host memory (train.py lazily loads + uploads one per view)
    gen = torch.Generator().manual_seed(11)
    gts_host = [torch.rand(3, H, W, generator=gen).pin_memory() for _ in range(4)]
    cams_host = cams
    for c in cams_host:
        for k in ("world_view_transform", "full_proj_transform", "camera_center"):
            setattr(c, k, getattr(c, k).pin_memory())
    cams_dev = [cam_to(c, device, non_blocking=False) for c in cams_host]
    gts_dev = [g.to(device) for g in gts_host]
    n_cams = len(cams_dev)
    state = {"R": 0, "vis": 0.0}

    def step(i, cam, gt):
        arena.zero()
        pkg = render(cam, scene, pipe, bg, 0.0, require_coord=False, require_depth=True, cam_no=cam.uid % 30,
                     iter=20000, num_down_emb_c=hyper.min_embeddings, num_down_emb_f=hyper.min_embeddings)
        loss = training_loss(pkg, gt, view=cam, reg=True)
        loss.backward()
        arena.allreduce(1.0 / world)
        state["pkg"] = pkg
        return loss

    def view_of(i):
        return (i * world + rank) % n_cams

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # BENCH_CLOCKS = thread (NVML from a thread of this process) | smi (nvidia-smi child, the recipe's line)
    sampler_cls = SmiClockSampler if os.environ.get("BENCH_CLOCKS", "thread") == "smi" else ClockSampler
    sampler = sampler_cls(local) if rank == 0 and os.environ.get("BENCH_CLOCKS") != "off" else None  # off: experiments only
    # untimed, before the W warm-up steps: every distinct view once, so that torch's caching allocator has seen the
    # largest per-view buffer sizes (a cudaMalloc inside the timed region is a multi-millisecond host stall)
    for v in range(n_cams):
        step(0, cams_dev[v], gts_dev[v % 4])
    for i in range(max(args.warmup, 3)):
        v = view_of(i)
        step(i, cams_dev[v], gts_dev[v % 4])
    barrier()
    grad_bytes = arena.nbytes()
    vis = state["pkg"]["visibility_filter"].float().mean().item()
    from ed3dgs_b200.rasterizer import _RasterizeGaussians  # noqa: F401
    R = int(state["pkg"]["render"].grad_fn.num_rendered) if hasattr(state["pkg"]["render"].grad_fn, "num_rendered") else 0

    # No cyclic-GC pass inside a timed region (a generation-2 sweep over the ctypes / tensor wrappers a step
    # creates shows up as a multi-millisecond bubble in a 16 ms step); reference counting still frees everything.
    import gc
    gc.collect()
    gc.disable()

    # ---- timed region 1: inputs resident in HBM ----
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    if sampler:
        sampler.mark_begin()
    mallocs0 = torch.cuda.memory_stats(device).get("num_device_alloc", 0)
    host_t = [time.perf_counter()]
    trace, trace_rows = bool(os.environ.get("BENCH_TRACE")), []
    ev0.record()
    for i in range(args.steps):
        v = view_of(i)
        step(i, cams_dev[v], gts_dev[v % 4])
        host_t.append(time.perf_counter())
        if trace:
            trace_rows.append((v, torch.cuda.memory_stats(device).get("num_device_alloc", 0),
                               torch.cuda.memory_reserved(device) >> 20))
    ev1.record()
    barrier()
    ms_total = ev0.elapsed_time(ev1)
    # host-side evidence for the region: per-step issue times (a step contains one blocking read of the instance
    # count, so host time tracks device time) and cudaMalloc calls made by the caching allocator inside it
    host_ms = sorted((b - a) * 1e3 for a, b in zip(host_t[:-1], host_t[1:]))
    host_info = {"step_issue_ms_median": host_ms[len(host_ms) // 2], "step_issue_ms_max": host_ms[-1],
                 "device_mallocs_in_region": torch.cuda.memory_stats(device).get("num_device_alloc", 0) - mallocs0}
    if os.environ.get("BENCH_TRACE"):
        print("view/mallocs/reserved_MiB:", " ".join(f"{a}/{b - mallocs0}/{c}" for a, b, c in trace_rows), file=sys.stderr)
        print("step issue ms:", " ".join(f"{(b - a) * 1e3:.1f}" for a, b in zip(host_t[:-1], host_t[1:])), file=sys.stderr)
    clocks = None
    if sampler:
        sampler.mark_end()
        clocks = sampler.stop()

    # ---- timed region 2 (e2e): host buffers, H2D of the step's inputs and D2H of the loss inside ----
    h2d = 3 * H * W * 4 + (16 + 16 + 3) * 4

    # The loop a training script with a prefetching loader runs: the inputs of step i + 1 are copied on a side
    # stream while step i computes, and the loss of every step is copied to pinned host memory without stalling
    # the launch of the next step. Every copy is issued (and completes) inside the timed region.
    copy_stream = torch.cuda.Stream(device=device)
    main_stream = torch.cuda.current_stream(device)
    loss_host = torch.zeros(args.steps + 3, dtype=torch.float32).pin_memory()

    # three device-side staging slots (allocated once: tensors created on a side stream every step made the
    # caching allocator defer their reuse and cudaMalloc again); slot s is rewritten only after the main stream
    # has passed the step that last read it
    NSLOT = 3
    slots = [{"gt": torch.empty(3, H, W, device=device),
              "mats": torch.empty(35, device=device),
              "done": None} for _ in range(NSLOT)]
    mats_host = [torch.cat([c.world_view_transform.reshape(-1), c.full_proj_transform.reshape(-1),
                            c.camera_center.reshape(-1)]).pin_memory() for c in cams_host]

    def prefetch(i):
        v = view_of(i)
        sl = slots[i % NSLOT]
        if sl["done"] is not None:
            copy_stream.wait_event(sl["done"])
        with torch.cuda.stream(copy_stream):
            sl["mats"].copy_(mats_host[v], non_blocking=True)
            sl["gt"].copy_(gts_host[v % 4], non_blocking=True)
            ready = torch.cuda.Event()
            ready.record(copy_stream)
        cam = NS(**vars(cams_host[v]))
        cam.world_view_transform = sl["mats"][0:16].view(4, 4)
        cam.full_proj_transform = sl["mats"][16:32].view(4, 4)
        cam.camera_center = sl["mats"][32:35]
        return cam, sl["gt"], ready, sl

    def e2e_loop(n, slot0):
        nxt = prefetch(0)
        for i in range(n):
            cam, gt, ready, sl = nxt
            if i + 1 < n:
                nxt = prefetch(i + 1)
            main_stream.wait_event(ready)
            loss = step(i, cam, gt)
            loss_host[slot0 + i].copy_(loss.detach(), non_blocking=True)   # device -> host read of the step's result
            sl["done"] = torch.cuda.Event()
            sl["done"].record(main_stream)

    e2e_loop(3, args.steps)   # untimed: lets the caching allocator settle on this region's allocation pattern
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    e2e_loop(args.steps, 0)
    e1.record()
    barrier()
    host_losses = [float(x) for x in loss_host[:args.steps]]
    ms_e2e = e0.elapsed_time(e1)
    gc.enable()

    # ---- multi-GPU breakdown (untimed for the headline): the collective alone, and steps without it ----
    multi = None
    if world > 1:
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        a0.record()
        for _ in range(10):
            arena.allreduce(1.0 / world)
        a1.record()
        barrier()
        c0, c1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        keep = arena.allreduce
        arena.allreduce = lambda *_a, **_k: keep(collective=False)
        c0.record()
        for i in range(10):
            v = view_of(i)
            step(i, cams_dev[v], gts_dev[v % 4])
        c1.record()
        barrier()
        arena.allreduce = keep
        mt = torch.tensor([a0.elapsed_time(a1) / 10, c0.elapsed_time(c1) / 10], dtype=torch.float64, device=device)
        mt_max = mt.clone()
        dist.all_reduce(mt_max, op=dist.ReduceOp.MAX)
        dist.all_reduce(mt, op=dist.ReduceOp.MIN)
        multi = {"allreduce_ms": float(mt_max[0]), "compute_only_ms_per_step_max_rank": float(mt_max[1]),
                 "compute_only_ms_per_step_min_rank": float(mt[1]),
                 "overlap": (transport if overlap else False), "overlap_note": overlap_note,
                 "sm_reserve": sm_reserve if overlap else 0,
                 "nccl_max_ctas": os.environ.get("NCCL_MAX_CTAS"),
                 "what": "allreduce_ms = the whole exchange alone (both buckets back to back); compute_only = the same "
                         "steps without any collective; with overlap on, bucket A (xyz, f_dc, f_rest, opacity, scaling, "
                         "rotation: 59 of 91 floats per Gaussian) runs next to the deformation network's backward, "
                         "which leaves sm_reserve SMs to NCCL; exposed share = ms_per_step - compute_only"}
        multi["exposed_ms"] = None  # filled below, once ms_per_step is reduced over ranks

    # ---- sustained leg (untimed for the headline): the same loop for >= BENCH_SUSTAINED_S seconds with the clock
    #      sampler running, so the line also says what a long training run holds once clocks settle ----
    sustained = None
    sus_s = float(os.environ.get("BENCH_SUSTAINED_S", "3"))
    if sus_s > 0:
        s_sampler = sampler_cls(local) if rank == 0 and os.environ.get("BENCH_CLOCKS") != "off" else None
        n_sus = max(args.steps, int(sus_s * 1e3 / max(ms_total / args.steps, 1e-3)) + 1)
        gc.disable()
        barrier()
        if s_sampler:
            s_sampler.mark_begin()
        s0, s1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s0.record()
        for i in range(n_sus):
            v = view_of(i)
            step(i, cams_dev[v], gts_dev[v % 4])
        s1.record()
        barrier()
        gc.enable()
        ms_sus = s0.elapsed_time(s1) / n_sus
        sc = None
        if s_sampler:
            s_sampler.mark_end()
            sc = s_sampler.stop()
        st = torch.tensor([ms_sus], dtype=torch.float64, device=device)
        if world > 1:
            dist.all_reduce(st, op=dist.ReduceOp.MAX)
        sustained = {"steps": n_sus, "seconds": float(st[0]) * n_sus / 1e3, "ms_per_step": float(st[0]),
                     "value": world * 1e3 / float(st[0]), "unit": UNIT, "clocks": sc}

    # ---- roofline pass: per-kernel CUDA-event times from the library's own scopes ----
    prof_steps = min(args.steps, 5)
    profile_enable(True)
    for i in range(prof_steps):
        v = view_of(i)
        step(i, cams_dev[v], gts_dev[v % 4])
    torch.cuda.synchronize()
    prof = profile_report()
    pairs = profile_pairs()
    profile_enable(False)

    t = torch.tensor([ms_total, ms_e2e], dtype=torch.float64, device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total, ms_e2e = float(t[0]), float(t[1])
    if multi is not None:
        multi["exposed_ms"] = ms_total / args.steps - multi["compute_only_ms_per_step_max_rank"]

    # ---- GPU reference leg (N = 1, rank 0, outside every product timing): BASELINE.md B1 + B2 ----
    gpu_ref = None
    if world == 1 and not args.no_gpu_reference:
        try:
            gpu_ref = gpu_reference_leg(cfg, scene, cams_dev, gts_dev, device, R, product_step=step)
        except Exception as ex:  # noqa: BLE001
            gpu_ref = {"unavailable": f"{type(ex).__name__}: {ex}"}
        torch.cuda.empty_cache()
    if rank != 0:
        return
    ms_step = ms_total / args.steps
    ms_step_e2e = ms_e2e / args.steps
    peaks = load_peaks()
    tf32_measured = measure_tf32_peak(device)
    tf32_peak = tf32_measured if tf32_measured else 0.5 * peaks["bf16_tflops"]
    tf32_src = ("cuBLAS tf32 measured in this run (torch.matmul fp32 8192^3, TF32 allowed, best of 10)" if tf32_measured
                else "half of the bf16 figure (tf32 probe failed)")
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    bits = 32 + max(1, math.ceil(math.log2(tiles + 1)))
    sort_passes = (bits + SORT_DIGIT_BITS - 1) // SORT_DIGIT_BITS
    work = algorithmic_work(cfg, hyper, R, vis)
    hbm_side = mlp_hbm_model(hyper, cfg["P"])
    breakdown = {k: ms / prof_steps for k, (ms, n) in prof.items()}
    launches = {k: n / prof_steps for k, (ms, n) in prof.items()}
    # DRAM traffic per step: NOT measured in this run (no profiler inside a timed bench). The figures are constants
    # read from an earlier `ncu --set full` capture of these kernels at C3 (TRAFFIC_FILE, same kernel sources); `roofline.traffic_source` says so on the line itself.
    traffic_source = None
    try:
        with open(os.path.join(ROOT, TRAFFIC_FILE)) as f:
            tj = json.load(f)
        measured = dict(tj["per_step_bytes"]) if args.workload == "C3" else {}
        if measured:
            if "tc_wgrad_ts_kernel" not in measured and "tc_wgrad_kernel" in measured:   # a capture with the older scope name
                measured["tc_wgrad_ts_kernel"] = measured.pop("tc_wgrad_kernel")          # (moved, not copied: summed once)
            traffic_source = (f"constant from {TRAFFIC_FILE} (ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum, "
                              f"an earlier capture of these kernels at C3), not measured in this run")
    except (OSError, KeyError, ValueError):
        measured = {}
    rooflines = {}
    for name, ms in breakdown.items():
        if name not in work or ms <= 0:
            continue
        bound, amount = work[name]
        nl = max(launches[name], 1)
        per_launch = amount / nl
        dur = ms / nl * 1e-3
        if bound == "hbm":
            ach, peak, unit = per_launch / dur / 1e9, peaks["hbm_gbs"], "GB/s"
        else:
            ach, peak, unit = per_launch / dur / 1e12, tf32_peak, "TFLOP/s"
        r = {"bound": bound, "achieved": ach, "peak": peak, "unit": unit, "frac": ach / peak,
             "traffic": (measured[name] / nl) if name in measured else None, "ms_per_step": ms,
             "launches_per_step": launches[name], "algorithmic_per_step": amount}
        if name in hbm_side:
            h = hbm_side[name]
            r["hbm_model"] = {"algorithmic_bytes_per_step": h["algorithmic"], "design_bytes_per_step": h["design"],
                              "design_gbs": h["design"] / (ms * 1e-3) / 1e9,
                              "design_frac_of_hbm_peak": h["design"] / (ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
                              "traffic_over_design": (measured[name] / h["design"]) if name in measured else None}
        rooflines[name] = r
    # section-level lines, the definitions of BASELINE.md section 5
    mlp_ms = sum(breakdown.get(k, 0.0) for k in MLP_GROUPS) + breakdown.get("fused_prepare", 0.0) + \
        breakdown.get("tc_prepare_b", 0.0) + breakdown.get("build_head_grads", 0.0)
    ras_ms = sum(breakdown.get(k, 0.0) for k in RASTER_GROUPS)
    sections = {}
    if mlp_ms > 0:
        m = mlp_shape(hyper)
        flop = 3 * work["fused_fwd_kernel"][1]
        mlp_traffic = sum(measured.get(k, 0) for k in MLP_GROUPS) or None
        sections["mlp"] = {"bound": "tensor", "ms_per_step": mlp_ms, "algorithmic_flop_per_step": flop,
                           "achieved": flop / (mlp_ms * 1e-3) / 1e12, "peak": tf32_peak, "unit": "TFLOP/s",
                           "frac": flop / (mlp_ms * 1e-3) / 1e12 / tf32_peak,
                           "frac_of_bf16_peak": flop / (mlp_ms * 1e-3) / 1e12 / peaks["bf16_tflops"],
                           "executed_over_algorithmic": 3.0, "launches_per_step": sum(
                               launches.get(k, 0) for k in MLP_GROUPS + ("fused_prepare", "tc_prepare_b", "build_head_grads")),
                           "hbm_algorithmic_bytes_per_step": sum(h["algorithmic"] for h in hbm_side.values()),
                           "hbm_traffic_bytes_per_step": mlp_traffic, "shape": m}
    if ras_ms > 0:
        P_, px_ = cfg["P"], W * H
        b_ras = P_ * (268 + 770 * vis) + 240 * R + 124 * px_
        sm_mhz = (clocks or {}).get("sm_max_mhz") or 1965.0
        fp32_peak = 148 * 128 * 2 * sm_mhz * 1e6
        e_f, e_b = (pairs[0] / prof_steps, pairs[1] / prof_steps) if pairs else (None, None)
        sec = {"bound": "hbm", "ms_per_step": ras_ms, "algorithmic_bytes_per_step": b_ras,
               "achieved": b_ras / (ras_ms * 1e-3) / 1e9, "peak": peaks["hbm_gbs"], "unit": "GB/s",
               "frac": b_ras / (ras_ms * 1e-3) / 1e9 / peaks["hbm_gbs"],
               "traffic_bytes_per_step": sum(measured.get(k, 0) for k in RASTER_GROUPS) or None,
               "fp32_alu": {"pairs_listed": R * 256, "pairs_walked_fwd": e_f, "pairs_walked_bwd": e_b,
                            "flop_per_pair": {"fwd": 30, "bwd": 110}, "peak_tflops": fp32_peak / 1e12,
                            "peak_how": "148 SMs x 128 FMA x 2 x max SM clock (derived, BASELINE.md 2)"}}
        blend_ms = breakdown.get("render_fwd", 0.0) + breakdown.get("blend_bwd", 0.0)
        if e_f is not None and blend_ms > 0:
            fl = 30 * e_f + 110 * e_b
            sec["fp32_alu"].update({"flop_per_step": fl, "blend_kernels_ms": blend_ms,
                                    "achieved_tflops": fl / (blend_ms * 1e-3) / 1e12,
                                    "frac": fl / (blend_ms * 1e-3) / fp32_peak,
                                    "frac_listed_pairs_model": 140 * R * 256 / (blend_ms * 1e-3) / fp32_peak})
        sections["rasterizer"] = sec
    dominant = max(rooflines, key=lambda k: rooflines[k]["ms_per_step"]) if rooflines else None
    roof = dict(rooflines[dominant], kernel=dominant, peak_source=(
        tf32_src if rooflines[dominant]["bound"] == "tensor" else "MEASURED_PEAKS.json" if peaks["source"] == "measured"
        else "B200_PROFILING.md fallback")) if dominant else None
    if roof:
        roof["traffic_source"] = traffic_source if roof.get("traffic") is not None else None
    if roof and roof["bound"] == "tensor" and roof.get("hbm_model", {}).get("design_frac_of_hbm_peak", 0) > 0.6:
        roof["note_hbm"] = ("this kernel is HBM-bound in practice: the bytes its design moves per launch (hbm_model."
                            "design_bytes_per_step, confirmed by ncu: traffic_over_design) over its time are %.0f %% of the "
                            "measured HBM copy peak; the tensor line is kept because SURVEY 8(d) defines the MLP roofline on "
                            "the tensor model" % (100 * roof["hbm_model"]["design_frac_of_hbm_peak"]))
    if roof and roof["bound"] == "tensor":
        roof["note"] = ("algorithmic flops = minimal formulation of SURVEY 8(d) for ONE pass (this kernel) over both "
                        "stages; the kernel executes 3 tf32 MMAs per algorithmic product (3xTF32, fp32-grade "
                        "parity), so frac <= 1/3 by construction")
    line = {
        "metric": METRIC, "value": world * 1e3 / ms_step, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32 (rasterizer fp32; MLP 3xTF32 on tcgen05, fp32-grade)", "data": "synthetic",
        "config": workload_config(args.workload, cfg, args.splat_scale),
        "workload_detail": workload_detail(
            cfg, R, world, vis,
            working_set="Gaussian records 64 MB + keys/lists %.0f MB + MLP activations %.1f GB per step"
                        % (R * 36 / 1e6, cfg["P"] * 7.0e3 / 1e9),
            grad_allreduce_bytes=grad_bytes if world > 1 else 0, multi_gpu=multi, host=host_info),
        "clocks": clocks,
        "e2e": {"value": world * 1e3 / ms_step_e2e, "unit": UNIT, "ms_per_step": ms_step_e2e,
                "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                "what": "render()+loss+backward through the reference-facing Python API; per step the camera "
                        "matrices and the target image are copied from pinned host memory (side stream, one step "
                        "ahead, like a prefetching loader) and the loss is copied back to pinned host memory"},
        "gpu_launches": launches_per_step(prof, prof_steps, sort_passes) * args.steps,
        "roofline": roof,
        "roofline_sections": sections,
        "roofline_all": rooflines,
        "sustained": sustained,
        "gpu_reference": gpu_ref,
        "loss_last": host_losses[-1] if host_losses else None,
    }
    if world == 1 and not args.no_cpu_baseline:
        try:
            ref = CpuReference(cfg, args.splat_scale)
            # three timed full-size steps after a warm-up step (about 7 s each at C3), capped at 30 s of timed work
            med, ts = ref.run(int(os.environ.get("BENCH_CPU_STEPS", "3")), 1, float(os.environ.get("BENCH_CPU_BUDGET_S", "30")))
            mean = sum(x[0] for x in ts) / len(ts)
            line["cpu_baseline"] = {"value": 1.0 / mean, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                                    "sample": ref.describe(med, ts, len(ts))}
        except Exception as ex:  # noqa: BLE001
            line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count(), "kind": "port",
                                    "sample": f"failed: {ex}"}
    print(json.dumps(line))


def gpu_reference_leg(cfg, scene, cams_dev, gts_dev, device, R_product, warm=3, iters=10, product_step=None):
    """BASELINE.md section 3, B1 + B2 on the same GPU and the same seeded inputs, outside the product's timed
    regions: the UNMODIFIED reference CUDA rasterizer (oracle/_ref, built from /root/reference by oracle/Makefile)
    wired through autograd like diff_gaussian_rasterization/__init__.py:46-174 wires it, the deformation network as
    the reference's own torch statements (oracle/deform_oracle.py: F.linear -> cuBLAS SGEMM, TF32 off) and the
    reference's torch losses. A reported baseline (the "3x" denominator); nothing here is on the product path."""
    import ctypes  # noqa: F401
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import deform_oracle
    import ref_binding
    from ed3dgs_b200 import synthetic
    from ed3dgs_b200.renderer import training_loss
    if not os.path.exists(ref_binding.REF_LIB):
        return {"unavailable": "oracle/_ref/libref_ed3dgs.so not built (needs /root/reference at build time)"}
    keep_tf32 = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    ref = ref_binding.RefRasterizer(sync=False)
    marks = {}

    def mark(name):
        e = torch.cuda.Event(enable_timing=True)
        e.record()
        marks[name] = e

    class RefRaster(torch.autograd.Function):
        @staticmethod
        def forward(ctx, means3D, sh, opac, scales, rots, rs, tongue):
            _R, out = ref.forward(rs, means3D.contiguous(), opac.contiguous(), tongue, shs=sh.contiguous(),
                                  scales=scales.contiguous(), rotations=rots.contiguous())
            return out["color"], out["depth"], out["mdepth"], out["alpha"], out["normal"]

        @staticmethod
        def backward(ctx, gc_, gd, gmd, ga, gn):
            mark("raster_bwd_begin")
            g = ref.backward(dict(color=gc_, depth=gd, mdepth=gmd, alpha=ga, normal=gn))
            mark("raster_bwd_end")
            return g["dL_dmeans3D"], g["dL_dsh"], g["dL_dopacity"], g["dL_dscales"], g["dL_drotations"], None, None

    params = dict(scene.named_parameters())
    sd = {k[len("_deformation."):]: v for k, v in params.items() if k.startswith("_deformation.")}
    P = cfg["P"]
    hyper = scene.hyper
    bg = torch.zeros(3, device=device)

    def ref_step(i):
        cam, gt = cams_dev[i % len(cams_dev)], gts_dev[i % len(gts_dev)]
        rs = synthetic.raster_settings(cam, cfg["sh_degree"], kernel_size=0.0, require_coord=False, require_depth=True, bg=bg)
        for p_ in scene.parameters():
            p_.grad = None
        mark("begin")
        time_emb = torch.full((P, 1), cam.time, device=device)
        shs = torch.cat((scene._features_dc, scene._features_rest), dim=1)
        out = deform_oracle.forward(sd, hyper, max(hyper.defor_depth, 1), hyper.max_embeddings, scene._xyz,
                                    scene._scaling, scene._rotation, scene._opacity, time_emb, cam.uid % 30,
                                    scene._embedding, shs, 20000, num_down_emb_c=hyper.min_embeddings,
                                    num_down_emb_f=hyper.min_embeddings)
        rots = torch.nn.functional.normalize(out[2])
        scales, opac = torch.exp(out[1]), torch.sigmoid(out[3])
        mark("mlp_fwd_end")
        color, depth, mdepth, alpha, normal = RefRaster.apply(out[0], out[4], opac, scales, rots, rs, scene.tongue_class)
        mark("raster_fwd_end")
        pkg = dict(render=color, expected_depth=depth, median_depth=mdepth, mask=alpha, normal=normal)
        training_loss(pkg, gt, view=cam, fused=False).backward()
        mark("end")

    try:
        for i in range(warm):
            ref_step(i)
        torch.cuda.synchronize()
        rows = []
        for i in range(iters):
            ref_step(i)
            torch.cuda.synchronize()
            m = marks
            rows.append(dict(total=m["begin"].elapsed_time(m["end"]),
                             mlp_fwd=m["begin"].elapsed_time(m["mlp_fwd_end"]),
                             raster_fwd=m["mlp_fwd_end"].elapsed_time(m["raster_fwd_end"]),
                             loss_fwd_bwd=m["raster_fwd_end"].elapsed_time(m["raster_bwd_begin"]),
                             raster_bwd=m["raster_bwd_begin"].elapsed_time(m["raster_bwd_end"]),
                             mlp_bwd=m["raster_bwd_end"].elapsed_time(m["end"])))
        med = {k: sorted(r[k] for r in rows)[len(rows) // 2] for k in rows[0]}
        same_views = None
        if product_step is not None:
            # The product on the IDENTICAL view list (views 0 .. iters-1, the ones timed above), split by the library's
            # own event scopes; both sides as means over that list, so the per-view spread of R cancels in the ratios.
            for p_ in scene.parameters():
                p_.grad = None
            profile_enable(True)
            profile_report()      # drop what earlier passes left in the log
            t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0.record()
            for i in range(iters):
                product_step(i, cams_dev[i % len(cams_dev)], gts_dev[i % len(gts_dev)])
            t1.record()
            torch.cuda.synchronize()
            pr = profile_report()
            profile_enable(False)
            mean = lambda k: sum(r[k] for r in rows) / len(rows)  # noqa: E731
            p_ras = sum(pr[k][0] for k in RASTER_GROUPS if k in pr) / iters
            p_mlp = sum(ms for k, (ms, _n) in pr.items()
                        if k in MLP_GROUPS or k in ("fused_prepare", "tc_prepare_b", "build_head_grads")) / iters
            p_step = t0.elapsed_time(t1) / iters     # (with the event scopes on: a few % above the untimed step)
            r_ras, r_mlp = mean("raster_fwd") + mean("raster_bwd"), mean("mlp_fwd") + mean("mlp_bwd")
            same_views = {"views": list(range(iters)), "reference_raster_ms": r_ras, "product_raster_ms": p_ras,
                          "raster_ratio": r_ras / p_ras if p_ras > 0 else None,
                          "reference_mlp_ms": r_mlp, "product_mlp_ms": p_mlp,
                          "mlp_ratio": r_mlp / p_mlp if p_mlp > 0 else None,
                          "reference_step_ms": mean("total"), "product_step_ms_profiled": p_step,
                          "step_ratio": mean("total") / p_step if p_step > 0 else None}
        return {"same_views": same_views, "what": "unmodified reference CUDA rasterizer (oracle/_ref) + the reference's torch fp32 deformation "
                        "network (cuBLAS SGEMM, TF32 off) + the reference's torch losses, same scene / views / loss, "
                        f"median of {iters} after {warm} warm-ups, CUDA events, outside the product's timed regions",
                "ms_per_step": med["total"], "value": 1e3 / med["total"], "unit": UNIT,
                "sections_ms": {k: v for k, v in med.items() if k != "total"},
                "mlp_ms": med["mlp_fwd"] + med["mlp_bwd"], "raster_ms": med["raster_fwd"] + med["raster_bwd"],
                "num_rendered_last": int(ref.R), "num_rendered_product": R_product}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = keep_tf32
        for p_ in scene.parameters():
            p_.grad = None


def main():
    args = parse()
    from ed3dgs_b200 import dp
    if args.impl == "reference":
        rank = int(os.environ.get("RANK", "0"))
        run_reference_arm(args, rank, int(os.environ.get("WORLD_SIZE", "1")))
        return
    if int(os.environ.get("WORLD_SIZE", "1")) > 1 and os.environ.get("BENCH_OVERLAP", "ce") in ("1", "nccl"):
        # the collective that runs next to the MLP backward gets exactly the SMs that backward leaves free
        os.environ.setdefault("NCCL_MAX_CTAS", os.environ.get("BENCH_SM_RESERVE", "8"))
    rank, local, world = dp.init_from_env("nccl")
    try:
        run_ours(args, rank, local, world)
    finally:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
