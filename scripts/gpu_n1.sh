#!/bin/bash
# One gpurun call: scripted-loop / variant-kernel bench lines, launch list and one full capture of the coupling kernel.
set -u
TAG=${1:-r01e}
OUT=gpurun_out
mkdir -p $OUT
timeout 600 python scripts/bench_scripted.py > $OUT/scripted_$TAG.json 2> $OUT/scripted_$TAG.err; echo "scripted rc=$?"; cat $OUT/scripted_$TAG.json; tail -5 $OUT/scripted_$TAG.err
timeout 300 python scripts/bench_multi_agent.py 1024 150 > $OUT/multi_agent_$TAG.json 2>> $OUT/scripted_$TAG.err; cat $OUT/multi_agent_$TAG.json
CMD="python scripts/bench_multi_agent.py 1024 20"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/launches_multi_agent_$TAG.csv $CMD > $OUT/ncu_ma_$TAG.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:coupling_kernel -s 6 -c 2 -f -o $OUT/prof_coupling_$TAG $CMD > $OUT/ncu_cp_$TAG.log 2>&1; echo "ncu full rc=$?"
ls -la $OUT | tail -8
