#!/bin/bash
# Multi-GPU sweep on one box (run through gpurun --gpus 8): C3 with / without the overlapped exchange at several NCCL CTA
# caps, C4 at 2/4/8 ranks, C5 at 8 ranks. One bench line per run under gpurun_out/<tag>/.
tag=${1:-dp}; out=gpurun_out/$tag; mkdir -p $out
run() {  # name nproc workload env...
  name=$1; np=$2; wl=$3; shift 3
  env BENCH_SUSTAINED_S=0 "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 \
    --master-port 29541 bench.py --gpus $np --workload $wl --steps 20 --warmup 3 > $out/$name.json 2> $out/$name.err
  python - <<PY
import json
try:
    d = json.loads(open("$out/$name.json").read().strip().splitlines()[-1])
    m = d["config"]["multi_gpu"] or {}
    print("$name", round(d["value"], 1), "views/s", round(d["ms_per_step"], 3), "ms; allreduce", round(m.get("allreduce_ms", 0), 3),
          "compute-only max", round(m.get("compute_only_ms_per_step_max_rank", 0), 3), "exposed", round(m.get("exposed_ms") or 0, 3))
except Exception as ex:
    print("$name failed", ex)
PY
}
n=${2:-8}
run c3_${n}gpu_single $n C3 BENCH_OVERLAP=0
run c3_${n}gpu_overlap_r4 $n C3 BENCH_OVERLAP=1 BENCH_SM_RESERVE=4 NCCL_MAX_CTAS=4
run c3_${n}gpu_overlap_r8 $n C3 BENCH_OVERLAP=1 BENCH_SM_RESERVE=8 NCCL_MAX_CTAS=8
run c3_${n}gpu_overlap_r16 $n C3 BENCH_OVERLAP=1 BENCH_SM_RESERVE=16 NCCL_MAX_CTAS=16
run c3_${n}gpu_overlap_r8_free $n C3 BENCH_OVERLAP=1 BENCH_SM_RESERVE=8 NCCL_MAX_CTAS=32
if [ "$n" = 8 ]; then
  run c3_4gpu_single 4 C3 BENCH_OVERLAP=0
  run c3_2gpu_single 2 C3 BENCH_OVERLAP=0
  run c4_2gpu_single 2 C4 BENCH_OVERLAP=0
  run c4_4gpu_single 4 C4 BENCH_OVERLAP=0
  run c4_8gpu_single 8 C4 BENCH_OVERLAP=0
  run c4_8gpu_overlap_r8 8 C4 BENCH_OVERLAP=1 BENCH_SM_RESERVE=8 NCCL_MAX_CTAS=8
  run c5_8gpu_single 8 C5 BENCH_OVERLAP=0
  run c5_8gpu_overlap_r8 8 C5 BENCH_OVERLAP=1 BENCH_SM_RESERVE=8 NCCL_MAX_CTAS=8
fi
ls $out
