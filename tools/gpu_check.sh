#!/bin/bash
# tests + smoke + bench summary on the GPU box (used through gpurun)
set -o pipefail
timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -5
timeout 120 python __graft_entry__.py --smoke 2>&1 | tail -4
OUT=${1:-gpurun_out/bench_convae.json}
timeout 600 python bench.py ${BENCH_ARGS} > $OUT 2> gpurun_out/bench_convae.err || tail -5 gpurun_out/bench_convae.err
python - <<PY
import json
d=json.load(open("$OUT"))
print({k:d[k] for k in ("value","ms_per_step","gpu_launches","clocks")}); print(d["e2e"]); print(d["roofline"]); print(d["cpu_baseline"])
for k,v in d["kernels"].items(): print("%-28s %6.1f launches %8.3f ms"%(k,v["launches_per_step"],v["ms_per_step"]))
PY
