#!/bin/bash
# usage: tools/gpu_quick.sh TAG [pytest-args]   -> tests (optional) + 512^3 bench, prints top kernels
TAG=$1; shift
if [ -n "$1" ]; then python -m pytest "$@" -m gpu -q -x 2>&1 | tail -4; fi
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_$TAG.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["step_hbm_frac"])
for k,v in list(d["kernels"].items())[:16]: print(k, v["ms_per_launch"], v["gbs"])
PY
