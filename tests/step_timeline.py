"""CPU-issue time vs GPU time of the phases of one training step (finds host-side bubbles)."""
import sys, time
from types import SimpleNamespace as NS
import torch
sys.path.insert(0, "..")
import conftest  # noqa
import bench
from ed3dgs_b200 import synthetic, dp
from ed3dgs_b200.renderer import render, training_loss

cfg = synthetic.CONFIGS["C3"]
dev = torch.device("cuda:0")
scene, cams = bench.build_scene(cfg, 0.3, dev)
arena = dp.GradSync(list(scene.parameters()))
cam = bench.cam_to(cams[3], dev, False)
gt = torch.rand(3, cfg["H"], cfg["W"], device=dev)
bg = torch.zeros(3, device=dev)
pipe = NS(debug=False)

def phase(name, fn, acc):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = fn()
    t1 = time.perf_counter()
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    acc.setdefault(name, []).append(((t1 - t0) * 1e3, (t2 - t0) * 1e3))
    return out

acc = {}
for it in range(6):
    phase("zero", arena.zero, acc)
    pkg = phase("render", lambda: render(cam, scene, pipe, bg, 0.0, require_coord=False, require_depth=True, cam_no=1,
                                         iter=20000, num_down_emb_c=30, num_down_emb_f=30), acc)
    loss = phase("loss", lambda: training_loss(pkg, gt, view=cam), acc)
    phase("backward", loss.backward, acc)
for k, v in acc.items():
    v = v[2:]
    print(f"{k:10s} cpu_issue {sum(a for a, b in v) / len(v):8.2f} ms   total(with sync) {sum(b for a, b in v) / len(v):8.2f} ms")
# unsynchronised loop for comparison
torch.cuda.synchronize(); t0 = time.perf_counter()
for it in range(5):
    arena.zero()
    pkg = render(cam, scene, pipe, bg, 0.0, require_coord=False, require_depth=True, cam_no=1, iter=20000,
                 num_down_emb_c=30, num_down_emb_f=30)
    training_loss(pkg, gt, view=cam).backward()
torch.cuda.synchronize(); print("free-running ms/step", (time.perf_counter() - t0) / 5 * 1e3)

def loop(n, sync_each):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for it in range(n):
        arena.zero()
        pkg = render(cam, scene, pipe, bg, 0.0, require_coord=False, require_depth=True, cam_no=1, iter=20000,
                     num_down_emb_c=30, num_down_emb_f=30)
        training_loss(pkg, gt, view=cam).backward()
        if sync_each:
            torch.cuda.synchronize()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / n * 1e3

import subprocess
print("free", loop(10, False), "sync_each", loop(10, True), "free", loop(10, False))
print(subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active", "--format=csv,noheader"], capture_output=True, text=True).stdout)
import gc
gc.disable()
print("free gc off", loop(10, False))
print("mem", torch.cuda.max_memory_allocated() / 1e9, torch.cuda.memory_reserved() / 1e9, torch.cuda.memory_stats()["num_alloc_retries"], torch.cuda.memory_stats().get("num_device_alloc", None))
