#!/bin/bash
# same-box sweep of the conv stack's utterances-per-pass: tools/ab_chunk.sh 7 14 30 60
for v in "$@"; do
  timeout 200 python bench.py --no-cpu-baseline --no-extras --steps ${STEPS:-30} --warmup 5 --chunk $v > gpurun_out/abc_$v.json 2> gpurun_out/abc.err || tail -3 gpurun_out/abc.err
  python - <<PY
import json
d=json.load(open("gpurun_out/abc_$v.json"))
k=d["kernels"]
print("chunk=$v: %.3f ms/step  e2e %.3f frac %.3f | "%(d["ms_per_step"],d["e2e"]["ms_per_step"],d["roofline"]["frac"])+" ".join("%s=%.0f"%(n.replace("tapgemm:","").replace("_kernel",""),1e3*x["ms_per_step"]) for n,x in k.items()))
PY
done
