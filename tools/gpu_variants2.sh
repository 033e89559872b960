#!/bin/bash
# usage: tools/gpu_variants2.sh TAG "pytest args" name1 name2 ...  -> optional tests, then the 512^3 bench with the default library and each variants/libmhh_b200_<name>.so
TAG=$1; shift; PT=$1; shift
mkdir -p gpurun_out
if [ -n "$PT" ]; then python -m pytest $PT -m gpu -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest exit $?"; tail -15 gpurun_out/pytest_$TAG.log; fi
run() {
  name=$1; shift
  env "$@" python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/bench_${TAG}_$name.json 2> gpurun_out/bench_${TAG}_$name.err; echo "== $name exit $?"; tail -3 gpurun_out/bench_${TAG}_$name.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/bench_${TAG}_$name.json").read().strip().splitlines()[-1])
print("$name", round(d["ms_per_step"],3), "ms/step", d["clocks"], "div", d["status"]["div"], "tke", d["status"]["tke"])
print("   ", " | ".join(f"{k.replace('_kernel','')} {v['ms_per_launch']}" for k,v in list(d["kernels"].items())[:13]))
PY
}
run default A=1
for v in "$@"; do run $v MHH_B200_LIB=$PWD/microhh_b200/variants/libmhh_b200_$v.so; done
