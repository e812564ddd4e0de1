#!/bin/bash
# Multi-GPU sweep on one box (run through gpurun --gpus 8): C3 / C4 / C5 with the single NCCL exchange at the end of the
# step and with the copy-engine exchange of bucket A next to the network's backward. One bench line per run under
# gpurun_out/<tag>/.   usage: bash tests/dp_sweep.sh <tag> [full|quick]
tag=${1:-dp}; out=gpurun_out/$tag; mkdir -p $out
run() {  # name nproc workload env...
  name=$1; np=$2; wl=$3; shift 3
  env BENCH_SUSTAINED_S=0 "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 \
    --master-port 29541 bench.py --gpus $np --workload $wl --steps 20 --warmup 3 > $out/$name.json 2> $out/$name.err
  python - <<PY
import json
try:
    d = json.loads(open("$out/$name.json").read().strip().splitlines()[-1])
    m = (d.get("workload_detail") or d["config"]).get("multi_gpu") or {}
    print("$name", round(d["value"], 1), "views/s", round(d["ms_per_step"], 3), "ms; allreduce", round(m.get("allreduce_ms", 0), 3),
          "compute-only max", round(m.get("compute_only_ms_per_step_max_rank", 0), 3), "exposed", round(m.get("exposed_ms") or 0, 3))
except Exception as ex:
    print("$name failed", ex)
PY
}
run c3_8gpu_single 8 C3 BENCH_OVERLAP=0
run c3_8gpu_ce 8 C3 BENCH_OVERLAP=ce
run c3_4gpu_ce 4 C3 BENCH_OVERLAP=ce
run c3_4gpu_single 4 C3 BENCH_OVERLAP=0
run c4_8gpu_ce 8 C4 BENCH_OVERLAP=ce
run c4_8gpu_single 8 C4 BENCH_OVERLAP=0
run c4_4gpu_ce 4 C4 BENCH_OVERLAP=ce
run c4_2gpu_ce 2 C4 BENCH_OVERLAP=ce
run c5_8gpu_ce 8 C5 BENCH_OVERLAP=ce
run c5_8gpu_single 8 C5 BENCH_OVERLAP=0
ls $out | wc -l
