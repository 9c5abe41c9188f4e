#!/bin/bash
# One gpurun call: config-5 sweep, multi-agent / scripted / variant bench lines with the current kernels.
set -u
TAG=${1:-r01h}
OUT=gpurun_out
mkdir -p $OUT
timeout 1500 python scripts/sweep_config5.py $OUT/config5_sweep_$TAG.json > $OUT/config5_$TAG.log 2>&1; echo "sweep rc=$?"; tail -3 $OUT/config5_$TAG.log
timeout 300 python scripts/bench_multi_agent.py 1024 150 > $OUT/multi_agent_$TAG.json 2> $OUT/ma_$TAG.err; cat $OUT/multi_agent_$TAG.json | cut -c1-300
timeout 600 python scripts/bench_scripted.py > $OUT/scripted_$TAG.json 2>> $OUT/ma_$TAG.err; cat $OUT/scripted_$TAG.json | cut -c1-600
