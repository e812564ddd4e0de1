"""Summarise the source page of an ncu report: hottest SASS lines by stall samples, per kernel id.
usage: python ncu_src_top.py report.ncu-rep [kernel_index] [top]"""
import csv, io, subprocess, sys
rep = sys.argv[1]
kid = int(sys.argv[2]) if len(sys.argv) > 2 else 0
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-kernel-base", "function"],
                     capture_output=True, text=True).stdout
# kernels are separated by a "Kernel Name" line
blocks, cur = [], None
for line in out.splitlines():
    if line.startswith('"Kernel Name"'):
        cur = [line]
        blocks.append(cur)
    elif cur is not None:
        cur.append(line)
blk = blocks[kid]
print(blk[0][:160])
rows = list(csv.reader(io.StringIO("\n".join(blk[1:]))))
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
samp = ix["# Samples"] if "# Samples" in ix else ix["Warp Stall Sampling (All Samples)"]
stall_cols = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = []
for n, r in enumerate(rows[1:]):
    try:
        s = int(r[samp])
    except (ValueError, IndexError):
        continue
    data.append((s, n, r))
tot = sum(d[0] for d in data)
print("total samples", tot, "lines", len(data))
for s, n, r in sorted(data, key=lambda d: -d[0])[:top]:
    st = sorted(((int(r[i] or 0), hdr[i][6:]) for i in stall_cols), reverse=True)[:2]
    print(f"{n:5d} {s:7d} {100.0*s/tot:5.1f}%  {r[ix['Source']].strip()[:70]:70s} {st}")
if len(sys.argv) > 5:
    lo, hi = int(sys.argv[4]), int(sys.argv[5])
    print("--- range")
    for s, n, r in sorted(data, key=lambda d: d[1]):
        if lo <= n <= hi and s >= 20:
            st = sorted(((int(r[i] or 0), hdr[i][6:]) for i in stall_cols), reverse=True)[:2]
            print(f"{n:5d} {s:7d}  {r[ix['Source']].strip()[:80]:80s} {st}")
if len(sys.argv) > 6:
    # region sums: boundaries given as comma list in argv[6]
    bounds = [int(x) for x in sys.argv[6].split(",")]
    print("--- regions")
    for a, b in zip(bounds[:-1], bounds[1:]):
        reg = [d for d in data if a <= d[1] < b]
        tot_r = sum(d[0] for d in reg)
        agg = {}
        for s, n, r in reg:
            for i in stall_cols:
                agg[hdr[i][6:]] = agg.get(hdr[i][6:], 0) + int(r[i] or 0)
        topst = sorted(agg.items(), key=lambda kv: -kv[1])[:5]
        print(f"[{a},{b}) samples={tot_r} {topst}")
