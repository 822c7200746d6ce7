#!/bin/bash
# same-box comparison of runtime variants selected by an environment variable: tools/ab_env.sh VAR v1 v2 ...
VAR=$1; shift
for rep in 1 2; do
  for v in "$@"; do
    env $VAR=$v timeout 200 python bench.py --no-cpu-baseline --no-extras --steps ${STEPS:-30} --warmup 5 > gpurun_out/abe_$v.json 2> gpurun_out/abe.err || tail -3 gpurun_out/abe.err
    python - <<PY
import json
d=json.load(open("gpurun_out/abe_$v.json"))
k=d["kernels"]
print("$VAR=$v rep$rep: %.3f ms/step  e2e %.3f frac %.3f | "%(d["ms_per_step"],d["e2e"]["ms_per_step"],d["roofline"]["frac"])+" ".join("%s=%.0f"%(n.replace("tapgemm:","").replace("_kernel",""),1e3*x["ms_per_step"]) for n,x in k.items()))
PY
  done
done
