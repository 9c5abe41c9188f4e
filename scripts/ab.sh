#!/bin/bash
# A/B of kernel variants on one box: scripts/ab.sh <tag> name1 name2 ...  (libmpcb_<name>.so built by scripts/build_variant.sh)
# extra bench arguments through AB_ARGS, e.g. AB_ARGS="--horizon 13"
TAG=$1; shift
mkdir -p gpurun_out
for rep in 1 2; do
for v in "$@"; do
  MPCB_LIB_PATH=$PWD/mpc_for_av_at_intersection_b200/libmpcb_$v.so timeout 300 python bench.py --steps 20 --warmup 5 --no-cpu-baseline $AB_ARGS > gpurun_out/ab_${TAG}_${v}_$rep.json 2> gpurun_out/ab_${TAG}_${v}_$rep.err
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/ab_${TAG}_${v}_$rep.json"))
    print(f"{'${v}':<16} rep $rep: {d['value']/1e6:.3f} M solves/s  {d['ms_per_step']:.4f} ms/step  p50 {d['p50_step_ms']:.4f}  iters {d['config']['mean_ipm_iters']:.3f}  failed {d['config']['failed_solves']}  frac {d['roofline']['frac']:.4f} e2e {d['e2e']['value']/1e6:.3f}")
except Exception as e:
    print("${v} rep $rep: FAILED", e)
PY
done
done
