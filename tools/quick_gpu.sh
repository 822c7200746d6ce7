#!/bin/bash
# quick GPU iteration: parity tests, a short bench with the per-kernel table, optionally a role trace (TRACE=1)
OUT=${1:-gpurun_out/quick.json}
timeout ${TEST_TIMEOUT:-150} python -m pytest tests -m gpu -q -x 2>&1 | tail -3
timeout ${BENCH_TIMEOUT:-120} python bench.py --no-cpu-baseline --steps ${STEPS:-30} --warmup 5 > $OUT 2> $OUT.err || tail -5 $OUT.err
python - <<PY
import json
d=json.load(open("$OUT"))
print("value %.0f  ms/step %.3f  e2e ms %.3f  frac %.3f  clocks %s"%(d["value"], d["ms_per_step"], d["e2e"]["ms_per_step"], d["roofline"]["frac"], d["clocks"]))
print(" ".join("%s=%.0f"%(n.replace("tapgemm:","").replace("_kernel",""),1e3*v["ms_per_step"]) for n,v in d["kernels"].items()))
PY
if [ -n "$TRACE" ]; then SPANON_TC_TRACE=1 timeout 120 python tools/prof_convae.py 60 1 2> $OUT.trace | tail -1; fi
