#!/bin/bash
# Round profile on the GPU box (through gpurun): plain run, then the ncu launch list and one --set full capture of a step.
# usage: tools/ncu_round.sh <tag>
TAG=${1:-r01}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain_$TAG.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_launch_$TAG.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'tapgemm_tc|fbank400|first_layer|pool_planar' -s 24 -c 12 \
    -o gpurun_out/prof_$TAG -f $CMD > gpurun_out/ncu_full_$TAG.log 2>&1
echo "rc=$?"; tail -3 gpurun_out/plain_$TAG.log; tail -3 gpurun_out/ncu_full_$TAG.log; ls -la gpurun_out/
