#!/bin/bash
# round-2 final single-GPU call: the no-allocation test alone in a fresh process, the full GPU suite, the bench, then the ncu launch list
TAG=${1:-r02x}
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/pytest_noalloc_$TAG.log 2>&1; echo "no_alloc alone exit $?"; tail -5 gpurun_out/pytest_noalloc_$TAG.log
python -m pytest tests -m gpu -q -x --durations=10 > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest exit $?"; tail -16 gpurun_out/pytest_$TAG.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; rc=$?; echo "bench exit $rc"; tail -3 gpurun_out/bench_$TAG.err
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_$TAG.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["value"], d["e2e"]["value"], d["roofline"], d.get("cpu_baseline"), d["clocks"], d["status"])
print(d.get("secondary"))
for k,v in list(d["kernels"].items())[:24]: print(k, v["ms_per_launch"], v["launches"], v["gbs"])
PY
if [ $rc -eq 0 ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-secondary > gpurun_out/ncu_launch_$TAG.log 2>&1; echo "ncu launches exit $?"
fi
