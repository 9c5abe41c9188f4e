#!/bin/bash
# Build a kernel variant of libmpcb.so next to the product library (A/B runs: MPCB_LIB_PATH=<file> python bench.py ...).
# usage: scripts/build_variant.sh <name> [extra nvcc flags]   ->  mpc_for_av_at_intersection_b200/libmpcb_<name>.so
set -e
NAME=$1; shift
D=$(dirname "$0")/../mpc_for_av_at_intersection_b200
nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC -cudart shared "$@" -o $D/libmpcb_$NAME.so $D/csrc/*.cu
echo built $D/libmpcb_$NAME.so
