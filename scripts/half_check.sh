#!/bin/bash
# One gpurun call for the two-problems-per-warp kernel: synccheck on a tiny run, the T = 10 / 13 parity tests, A/B against the
# lane-per-stage kernel (MPCB_HALF_WARP=0).  Every step under its own short timeout (a new kernel may hang).
set -u
OUT=gpurun_out; mkdir -p $OUT
timeout 120 python __graft_entry__.py smoke > $OUT/half_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $OUT/half_smoke.log
timeout 600 python -m pytest tests -m gpu -x -q -s > $OUT/half_pytest.log 2>&1; echo "pytest rc=$?"; grep "passed\|failed\|Error\|error" $OUT/half_pytest.log | tail -8
for rep in 1 2; do
for T in 13 10; do
for hw in 0 1; do
  MPCB_HALF_WARP=$hw timeout 200 python bench.py --horizon $T --steps 20 --warmup 5 --no-cpu-baseline > $OUT/half_T${T}_hw${hw}_$rep.json 2> $OUT/half_T${T}_hw${hw}_$rep.err
  python - <<PY
import json
try:
    d = json.load(open("$OUT/half_T${T}_hw${hw}_$rep.json"))
    print(f"T=$T half=$hw rep $rep: {d['value']/1e6:.3f} M solves/s  {d['ms_per_step']:.4f} ms/step  iters {d['config']['mean_ipm_iters']:.3f}  failed {d['config']['failed_solves']}  e2e {d['e2e']['value']/1e6:.3f}")
except Exception as e:
    print("T=$T half=$hw rep $rep: FAILED", e)
PY
done; done; done
MPCB_HALF_WARP=1 timeout 200 python bench.py --horizon 13 --batch 65536 --steps 20 --warmup 5 --no-cpu-baseline > $OUT/half_T13_b65536.json 2>/dev/null; cut -c1-200 $OUT/half_T13_b65536.json
MPCB_HALF_WARP=0 timeout 200 python bench.py --horizon 13 --batch 65536 --steps 20 --warmup 5 --no-cpu-baseline > $OUT/half_T13_b65536_hw0.json 2>/dev/null; cut -c1-200 $OUT/half_T13_b65536_hw0.json
