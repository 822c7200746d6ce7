#!/bin/bash
# same-box comparison of several builds of libspanon.so: tools/ab_multi.sh tools/ab/X.so tools/ab/Y.so ... (2 reps, alternating)
LIB=speech_anonymization_b200/libspanon.so
cp $LIB /tmp/keep.so
for rep in 1 2; do
  for f in "$@"; do
    cp $f $LIB
    v=$(basename $f .so)
    timeout 300 python bench.py --no-cpu-baseline --no-extras --steps ${STEPS:-20} --warmup 5 > gpurun_out/abm_$v.json 2> gpurun_out/abm.err || tail -3 gpurun_out/abm.err
    python - <<PY
import json
d=json.load(open("gpurun_out/abm_$v.json"))
k=d["kernels"]
tg=[(n,x["ms_per_step"]) for n,x in k.items() if "tapgemm" in n or "first_layer" in n]
print("$v rep$rep: %.3f ms/step conv %.0f | "%(d["ms_per_step"],1e3*sum(x for _,x in tg))+" ".join("%s=%.0f"%(n.replace("tapgemm:","").replace("_kernel",""),1e3*x) for n,x in tg))
PY
  done
done
cp /tmp/keep.so $LIB
