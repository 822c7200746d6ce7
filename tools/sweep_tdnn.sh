#!/bin/bash
# sweep of the TDNN dataflow launch's split / group parameters (same box): tools/sweep_tdnn.sh "68,40,40:8" "84,32,32:1" ...
for cfg in "$@"; do
  split=${cfg%%:*}; group=${cfg##*:}
  SPANON_TDNN_SPLIT=$split SPANON_TDNN_GROUP=$group timeout 200 python bench.py --no-cpu-baseline --no-extras --steps 10 --warmup 3 > gpurun_out/sw.json 2> gpurun_out/sw.err || tail -3 gpurun_out/sw.err
  python - <<PY
import json
d=json.load(open("gpurun_out/sw.json"))
k=d["kernels"]
print("$cfg: %.3f ms/step | "%(d["ms_per_step"])+" ".join("%s=%.0f"%(n.replace("tapgemm:","").replace("_kernel",""),1e3*x["ms_per_step"]) for n,x in k.items() if "tdnn" in n or "pool" in n))
PY
done
