#!/bin/bash
# Compare kernel tuning variants (MPCB_TUNE_VARIANTS build): parity tests on the default, bench on each.
set -u
TAG=${1:-tune}
OUT=gpurun_out
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/pytest_gpu_$TAG.log
for v in ${VARIANTS:-0 1 2 3}; do
  for b in ${BATCHES:-4096 65536}; do
    MPCB_VARIANT=$v timeout 300 python bench.py --no-cpu-baseline --steps ${STEPS:-40} --batch $b > $OUT/bench_${TAG}_v${v}_b${b}.json 2> $OUT/bench_${TAG}_v${v}_b${b}.err
    python - <<PY
import json
try:
    d=json.load(open("$OUT/bench_${TAG}_v${v}_b${b}.json"))
    print("variant $v batch $b: %.3f M solves/s  ms/step %.3f  e2e %.3f M  frac %.4f  iters %.2f fail %d" % (d["value"]/1e6, d["ms_per_step"], d["e2e"]["value"]/1e6, d["roofline"]["frac"], d["config"]["mean_ipm_iters"], d["config"]["failed_solves"]))
except Exception as e:
    print("variant $v batch $b: FAILED", e); print(open("$OUT/bench_${TAG}_v${v}_b${b}.err").read()[-800:])
PY
  done
done
