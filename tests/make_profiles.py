"""Turn one gpurun_out/<tag>/ directory (written by tests/gpu_round.sh) into the tracked files under profiles/:
  profiles/<name>_bench.json          the bench line as printed
  profiles/<name>_launches_step.csv   one steady-state step of the ncu launch list of the same bench command
  profiles/<name>_launches_step.md    the same step aggregated per kernel, with each kernel's share of the step
  profiles/<name>_ncu.md              one row per `ncu --set full` launch (MLP and rasterizer captures)
  profiles/<round>_traffic.json       DRAM bytes per step per kernel group (bench.py reads it for `roofline.traffic`)
usage: python tests/make_profiles.py gpurun_out/<tag> <name>      (run in the dev container; needs ncu for -i)"""
import collections
import csv
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src, name = sys.argv[1], sys.argv[2]
prof = os.path.join(ROOT, "profiles")


def short(n):
    n = re.sub(r"\(.*", "", n)
    return n.replace("ed3dgs::", "").replace("void ", "")[-70:].strip()


# ---- bench line ----
bj = os.path.join(src, "bench.json")
if os.path.exists(bj):
    line = open(bj).read().strip().splitlines()[-1]
    json.loads(line)
    open(os.path.join(prof, f"{name}_bench.json"), "w").write(line + "\n")

# ---- launch list: one step ----
lc = os.path.join(src, "launches_bench.csv")
step_share = {}
if os.path.exists(lc):
    rows = list(csv.reader(open(lc)))
    hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
    hdr = rows[hi]
    ix = {h: i for i, h in enumerate(hdr)}
    data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
    names = [r[ix["Kernel Name"]] for r in data]
    starts = [i for i, n in enumerate(names) if "temb_forward" in n]
    a, b = starts[-3], starts[-2]
    with open(os.path.join(prof, f"{name}_launches_step.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["ID", "Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum [ns]"])
        for r in data[a:b]:
            w.writerow([r[ix["ID"]], r[ix["Kernel Name"]][:160], r[ix["Block Size"]], r[ix["Grid Size"]], r[ix["Metric Value"]]])
    agg = collections.OrderedDict()
    for r in data[a:b]:
        k = short(r[ix["Kernel Name"]])
        e = agg.setdefault(k, [0.0, 0, "ed3dgs::" in r[ix["Kernel Name"]]])
        e[0] += float(r[ix["Metric Value"]].replace(",", ""))
        e[1] += 1
    tot = sum(v[0] for v in agg.values())
    ours = sum(v[1] for v in agg.values() if v[2])
    with open(os.path.join(prof, f"{name}_launches_step.md"), "w") as f:
        f.write(f"One steady-state step of `ncu --metrics gpu__time_duration.sum --clock-control none python bench.py --steps 3 "
                f"--warmup 3` (cold-cache, serialised): {b - a} launches ({ours} of this library, {b - a - ours} torch), "
                f"{tot / 1e6:.3f} ms of kernel time.\n\n| kernel | launches | us | share |\n|---|---|---|---|\n")
        for k, (v, c, _) in sorted(agg.items(), key=lambda x: -x[1][0]):
            f.write(f"| `{k}` | {c} | {v / 1e3:.1f} | {100 * v / tot:.1f} % |\n")
            step_share[k] = v / tot

# ---- ncu --set full captures ----
WANT = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "rd"), ("dram__bytes_write.sum", "wr"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem_lsu%"),
        ("l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem_tc%"),
        ("lts__t_sector_hit_rate.pct", "L2hit%"), ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"), ("launch__registers_per_thread", "regs"),
        ("smsp__inst_executed.sum", "inst"), ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%")]
traffic = collections.OrderedDict()
md = []
for rep, what in (("mlp.ncu-rep", "tests/profile_deform.py 1000000 3 (deformation network fwd+bwd, both stages, P = 1 M)"),
                  ("raster.ncu-rep", "tests/profile_raster.py mine C3 0.3 (rasterizer fwd+bwd, C3)")):
    path = os.path.join(src, rep)
    if not os.path.exists(path):
        continue
    raw = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    ix = {h: i for i, h in enumerate(hdr)}
    cols = [(ix[m], n, units[ix[m]]) for m, n in WANT if m in ix]
    md.append(f"\n### `ncu --set full --clock-control none --import-source on` of {what}\n")
    md.append("| # | kernel | grid | " + " | ".join(f"{n} [{u}]" for _, n, u in cols) + " |")
    md.append("|---|---|---|" + "---|" * len(cols))
    per_kernel = collections.OrderedDict()
    for r in rows[2:]:
        if len(r) < len(hdr):
            continue
        k = short(r[ix["Kernel Name"]])
        vals = []
        for i, n, u in cols:
            v = r[i].replace(",", "")
            try:
                fv = float(v)
                vals.append(f"{fv:.3g}" if abs(fv) < 1e4 else f"{fv:.4g}")
            except ValueError:
                vals.append(v)
        md.append(f"| {r[ix['ID']]} | `{k}` | {r[ix['Grid Size']]} | " + " | ".join(vals) + " |")

        def num(m):
            v, u = float(r[ix[m]].replace(",", "")), units[ix[m]]
            return v * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}.get(u, 1.0)
        per_kernel.setdefault(k, []).append(num("dram__bytes_read.sum") + num("dram__bytes_write.sum"))
    for k, v in per_kernel.items():
        traffic[k] = {"launches_captured": len(v), "bytes_per_launch": sum(v) / len(v)}
if md:
    with open(os.path.join(prof, f"{name}_ncu.md"), "w") as f:
        f.write(f"# ncu --set full summaries ({name}); per-launch times here are cold-cache and serialised\n" + "\n".join(md) + "\n")

# ---- traffic per step: bytes per launch x launches per step of the bench line ----
if traffic and os.path.exists(bj):
    d = json.loads(open(bj).read().strip().splitlines()[-1])
    per_step = {}
    for k, t in traffic.items():
        base = re.sub(r"<.*", "", k)
        group = {"radix_onesweep": "radix_sort_pairs", "radix_hist_all": "radix_sort_pairs",
                 }.get(base, base)
        if group == "tc_wgrad_ts_kernel" and group not in d["roofline_all"]:
            group = "tc_wgrad_kernel"      # bench lines before round 3 carried the older scope name
        if group in d["roofline_all"]:
            nl = d["roofline_all"][group]["launches_per_step"]
            if group == "radix_sort_pairs":
                n_here = 6 if base == "radix_onesweep" else 1
                per_step[group] = per_step.get(group, 0.0) + t["bytes_per_launch"] * n_here
            else:
                per_step[group] = t["bytes_per_launch"] * nl
    out = {"source": f"ncu --set full --clock-control none ({name}): dram__bytes_read.sum + dram__bytes_write.sum per launch "
                     f"(profiles/{name}_ncu.md) x launches per step of the bench line (profiles/{name}_bench.json)",
           "per_step_bytes": per_step, "per_launch": traffic}
    json.dump(out, open(os.path.join(prof, name.split("_")[0] + "_traffic.json"), "w"), indent=1)
    print(json.dumps(per_step, indent=1))
print("wrote profiles for", name)
