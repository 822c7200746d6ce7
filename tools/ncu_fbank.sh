#!/bin/bash
# ncu --set full of the two Fbank kernels (one launch each) on the bench shape; reports land in gpurun_out/
export PYTHONPATH=$PWD
for v in 0 1; do
  SPANON_FBANK_WARP=$v ncu --set full --clock-control none --import-source on -k regex:fbank400 -s 2 -c 1 -f -o gpurun_out/fbank_warp$v python tools/fb_once.py > gpurun_out/ncu_fbank$v.log 2>&1 || tail -5 gpurun_out/ncu_fbank$v.log
done
ls -la gpurun_out/*.ncu-rep
