#!/bin/bash
# One ncu --set full capture of the kernels matching a demangled-name regex in a short bench run.
# usage: tools/ncu_one.sh <out-tag> <regex> [skip] [count]
TAG=$1; RE=$2; SKIP=${3:-2}; CNT=${4:-1}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"$RE" -s $SKIP -c $CNT \
    -o gpurun_out/one_$TAG -f $CMD > gpurun_out/ncu_one_$TAG.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/ncu_one_$TAG.log; ls -la gpurun_out/one_$TAG.ncu-rep
