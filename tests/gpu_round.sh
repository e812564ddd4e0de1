#!/bin/bash
# One GPU-box round (run through gpurun): GPU tests, bench line, ncu launch list of the same bench command,
# ncu --set full of the dominant kernels. Everything lands under gpurun_out/<tag>/.
#   usage: bash tests/gpu_round.sh <tag> [tests|notests] [ncu|noncu]
tag=${1:-round}; want_tests=${2:-tests}; want_ncu=${3:-ncu}
out=gpurun_out/$tag; mkdir -p $out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/smi.txt 2>&1
if [ "$want_tests" = tests ]; then
  timeout 2400 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu.log
  tail -3 $out/pytest_gpu.log
fi
timeout 900 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
tail -c 600 $out/bench.err
python - <<EOF
import json
try:
    d = json.loads(open("$out/bench.json").read().strip().splitlines()[-1])
    print("ms/step", d["ms_per_step"], "e2e", d["e2e"]["ms_per_step"], "sustained", (d.get("sustained") or {}).get("ms_per_step"))
    print({k: round(v["ms_per_step"], 3) for k, v in d["roofline_all"].items()})
    print("gpu_reference", json.dumps(d.get("gpu_reference"))[:600])
except Exception as ex:
    print("bench parse failed", ex)
EOF
if [ "$want_ncu" = ncu ]; then
  BENCH_SUSTAINED_S=0 BENCH_CLOCKS=off timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv \
    --log-file $out/launches_bench.csv python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-gpu-reference > $out/ncu_bench.log 2>&1
  echo "ncu launch list exit $?"
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'fused_fwd_kernel|fused_bwd_kernel|tc_wgrad|tc_rows_kernel' \
    -s 6 -c 8 -o $out/mlp -f python tests/profile_deform.py 1000000 3 > $out/ncu_mlp.log 2>&1
  echo "ncu mlp exit $?"
  timeout 900 ncu --set full --clock-control none -k regex:'blend_bwd|render_fwd|gaussian_bwd|preprocess_fwd|onesweep|radix|emit_keys' \
    -s 13 -c 13 -o $out/raster -f python tests/profile_raster.py mine C3 0.3 > $out/ncu_raster.log 2>&1
  echo "ncu raster exit $?"
fi
for wl in C4 C5; do
  BENCH_SUSTAINED_S=0 timeout 400 python bench.py --workload $wl --steps 20 --warmup 3 --no-cpu-baseline > $out/bench_${wl}_1gpu.json 2> $out/bench_${wl}_1gpu.err
  echo "bench $wl exit $?"
done
du -sh $out; ls -la $out
