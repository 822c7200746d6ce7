#!/bin/bash
# bench.py over conv-stack chunk sizes (utterances per chunk); prints step time and the per-kernel table
for c in ${CHUNKS:-6 10 15 20 30 60}; do
  timeout 300 python bench.py --no-cpu-baseline --chunk $c --steps 8 --warmup 3 > gpurun_out/sweep_c$c.json 2> gpurun_out/sweep.err || tail -3 gpurun_out/sweep.err
  python - <<PY
import json
d=json.load(open("gpurun_out/sweep_c$c.json"))
k=d["kernels"]
print("chunk %3d: %.3f ms/step  e2e %.3f  | "%($c,d["ms_per_step"],d["e2e"]["ms_per_step"])+" ".join("%s=%.0f"%(n.replace("tapgemm:","").replace("_kernel",""),1e3*v["ms_per_step"]) for n,v in k.items()))
PY
done
