#!/bin/bash
# One gpurun call: GPU parity tests, smoke, bench lines (driver protocol), ncu launch list + one full capture of the step kernel.
# usage: gpurun --timeout 1800 -- 'bash scripts/gpu_check.sh [tag]'
set -u
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/gpu_$TAG.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q -s > $OUT/pytest_gpu_$TAG.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/pytest_gpu_$TAG.log
grep "passed\|failed" $OUT/pytest_gpu_$TAG.log | tail -3
timeout 300 python __graft_entry__.py smoke > $OUT/smoke_$TAG.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/smoke_$TAG.log
timeout 600 python bench.py --steps 20 --warmup 5 > $OUT/bench_$TAG.json 2> $OUT/bench_$TAG.err; echo "bench rc=$?"; cut -c1-600 $OUT/bench_$TAG.json; tail -3 $OUT/bench_$TAG.err
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $OUT/bench_ref_$TAG.json 2>> $OUT/bench_$TAG.err; cut -c1-300 $OUT/bench_ref_$TAG.json
timeout 300 python bench.py --batch 65536 --steps 20 --warmup 5 --no-cpu-baseline > $OUT/bench_b65536_$TAG.json 2>> $OUT/bench_$TAG.err; cut -c1-300 $OUT/bench_b65536_$TAG.json
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
timeout 300 $CMD > $OUT/plain_$TAG.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $OUT/launches_$TAG.csv $CMD > $OUT/ncu_launches_$TAG.log 2>&1
echo "ncu launches rc=$?"
timeout 300 $CMD > $OUT/plain2_$TAG.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mpc_step_kernel -s 8 -c 2 -f -o $OUT/prof_step_$TAG $CMD > $OUT/ncu_full_$TAG.log 2>&1
echo "ncu full rc=$?"
ls -la $OUT | tail -20
