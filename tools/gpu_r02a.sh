#!/bin/bash
# round-2 first GPU call: full GPU test suite (new large-grid parity tests), then the 512^3 bench with the gabls1 config
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x --durations=15 > gpurun_out/pytest_r02a.log 2>&1; echo "pytest exit $?"; tail -30 gpurun_out/pytest_r02a.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_r02a.json 2> gpurun_out/bench_r02a.err; echo "bench exit $?"; tail -5 gpurun_out/bench_r02a.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_r02a.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["value"], d["e2e"]["value"], d["roofline"], d.get("cpu_baseline"), d["clocks"], d["status"])
print(d.get("secondary"))
for k,v in list(d["kernels"].items())[:24]: print(k, v["ms_per_launch"], v["launches"], v["gbs"])
PY
