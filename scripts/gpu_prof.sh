#!/bin/bash
# ncu --set full on the step kernel for the given MPCB_VARIANT values (tuning builds) -> gpurun_out/prof_<tag>_v<k>.ncu-rep
set -u
TAG=${1:-p}
OUT=gpurun_out; mkdir -p $OUT
for v in ${VARIANTS:-0}; do
  CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --batch ${BATCH:-4096}"
  MPCB_VARIANT=$v timeout 300 $CMD > $OUT/plain_${TAG}_v$v.log 2>&1 &&
  MPCB_VARIANT=$v timeout 900 ncu --set full --clock-control none --import-source on -k regex:mpc_step_kernel -s 4 -c 1 -f -o $OUT/prof_${TAG}_v$v $CMD > $OUT/ncu_${TAG}_v$v.log 2>&1
  echo "variant $v ncu rc=$?"
done
