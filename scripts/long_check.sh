#!/bin/bash
# long-horizon kernel: parity tests at T = 40 / 80, then throughput at B = 4096 for thread-count variants
TAG=${1:-long}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity_step.py tests/test_gpu_bench_workloads.py -m gpu -q -s -k "40 or 80 or hard" > gpurun_out/long_tests_$TAG.log 2>&1
echo "tests rc=$?"; grep "T=\|hard cases\|passed\|failed\|Error" gpurun_out/long_tests_$TAG.log | cut -c1-220 | tail -12
run() {  # T threads steps
  MPCB_LONG_THREADS=$2 timeout 300 python bench.py --horizon $1 --batch ${4:-4096} --steps $3 --no-cpu-baseline > gpurun_out/long_T$1_nt$2_$TAG.json 2>gpurun_out/long_T$1_nt$2_$TAG.err
  python -c "
import json; d=json.load(open('gpurun_out/long_T$1_nt$2_$TAG.json')); print('T=$1 threads $2 B=${4:-4096}: %.1f K solves/s  %.3f ms/step  iters %.2f failed %d frac %.4f' % (d['value']/1e3, d['ms_per_step'], d['config']['mean_ipm_iters'], d['config']['failed_solves'], d['roofline']['frac']))" 2>&1 | tail -1
}
run 40 64 10; run 40 128 10; run 80 128 5; run 80 256 5
