#!/bin/bash
# A/B two builds of libspanon.so on the SAME box: tools/ab/A.so vs tools/ab/B.so, alternating runs.
# (copies each over the package's library; the last one stays in place)
LIB=speech_anonymization_b200/libspanon.so
for rep in 1 2; do
  for v in A B; do
    cp tools/ab/$v.so $LIB
    timeout 300 python bench.py --no-cpu-baseline --steps ${STEPS:-30} --warmup 5 > gpurun_out/ab_$v$rep.json 2> gpurun_out/ab.err || tail -3 gpurun_out/ab.err
    python - <<PY
import json
d=json.load(open("gpurun_out/ab_$v$rep.json"))
k=d["kernels"]
print("$v$rep: %.3f ms/step  e2e %.3f | "%(d["ms_per_step"],d["e2e"]["ms_per_step"])+" ".join("%s=%.0f"%(n.replace("tapgemm:","").replace("_kernel",""),1e3*v["ms_per_step"]) for n,v in k.items() if "tapgemm" in n))
PY
  done
done
