"""Summarise `ncu -i rep --page raw --csv` output: one markdown row per launch with the metrics the
roofline discussion uses. usage: python tests/ncu_raw_summary.py raw.csv"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
want = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "rd"), ("dram__bytes_write.sum", "wr"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor%"),
        ("sm__inst_executed_pipe_tensor_op_utcmma.avg.pct_of_peak_sustained_active", "utcmma%"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem_lsu%"),
        ("l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem_tc%"),
        ("lts__t_sector_hit_rate.pct", "L2hit%"), ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm%"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram%"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occ%"),
        ("launch__registers_per_thread", "regs"), ("smsp__inst_executed.sum", "inst"),
        ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "fma%"),
        ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu%"),
        ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue%")]
units = rows[1]
cols = [(ix[m], n, units[ix[m]]) for m, n in want if m in ix]
print("| # | kernel | grid | " + " | ".join(f"{n} [{u}]" for _, n, u in cols) + " |")
print("|---|---|---|" + "---|" * len(cols))
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    name = r[ix["Kernel Name"]].split("(")[0][-40:]
    vals = []
    for i, n, u in cols:
        v = r[i].replace(",", "")
        try:
            f = float(v)
            vals.append(f"{f:.3g}" if abs(f) < 1e4 else f"{f:.4g}")
        except ValueError:
            vals.append(v)
    print(f"| {r[ix['ID']]} | {name} | {r[ix['Grid Size']]} | " + " | ".join(vals) + " |")
