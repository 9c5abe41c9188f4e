#!/bin/bash
# usage: scripts/ncu_phase.sh <report.ncu-rep> <mangled kernel substring> [problems] [source file with '// [phase: ..]' markers]  -> per-phase instruction / sample table
set -e
R=$1; K=${2:-mpc_step_kernelILi20ELi1ELi8}; NP=${3:-4096}; SRC=${4:-mpcb_step.cuh}
W=$(mktemp -d)
ncu -i $R --page source --csv --print-source sass > $W/sass.csv 2>/dev/null
ncu -i $R --page raw --csv > $W/raw.csv
(cd $W && cuobjdump -xelf all ${NCU_PHASE_LIB:-/root/repo/mpc_for_av_at_intersection_b200/libmpcb.so} >/dev/null 2>&1 && nvdisasm -g mpcb_api.sm_100a.cubin > disall.txt)
python - $W/disall.txt $K > $W/dis.txt <<'PY'
import sys
L=open(sys.argv[1]).read().split('\n')
st=[i for i,l in enumerate(L) if l.startswith('.text.') and sys.argv[2] in l][0]
en=[i for i,l in enumerate(L) if l.startswith('.text.') and i>st]
print('\n'.join(L[st-8:(en[0]-8 if en else len(L))]))
PY
python scripts/ncu_phase.py $W/dis.txt $W/sass.csv $NP $SRC > $W/table.txt   # set -e: a failure prints nothing (no traceback posing as a table)
cat $W/table.txt
