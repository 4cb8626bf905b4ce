#!/bin/bash
# ncu full capture of the tcgen05 kernels alone.  Usage: bash tools/gpu_prof.sh <tag> <fwd|train> [kernel regex] [count]
TAG=$1; MODE=${2:-train}; RE=${3:-mlp_tc}; CNT=${4:-3}
OUT=gpurun_out; mkdir -p $OUT
timeout 300 python tools/prof_mlp.py $MODE 2 > $OUT/prof_plain_$TAG.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:$RE -c $CNT -o $OUT/prof_$TAG -f \
  python tools/prof_mlp.py $MODE 2 > $OUT/prof_ncu_$TAG.log 2>&1; echo "ncu exit $?"; tail -3 $OUT/prof_ncu_$TAG.log
