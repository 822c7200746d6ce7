#!/bin/bash
# ncu --set full of the Fbank kernel forms (one launch each) on the bench shape; reports land in gpurun_out/
export PYTHONPATH=$PWD
for v in ${@:-rr stockham warp}; do
  SPANON_FBANK=$v ncu --set full --clock-control none --import-source on -k regex:fbank400 -s 2 -c 1 -f -o gpurun_out/fbank_$v python tools/fb_once.py > gpurun_out/ncu_fbank_$v.log 2>&1 || tail -5 gpurun_out/ncu_fbank_$v.log
done
ls -la gpurun_out/fbank_*.ncu-rep
