#!/bin/bash
# usage: tools/gpu_ab.sh TAG "ENV=a ENV=b ..."  -> the 512^3 bench once per environment setting on the same box (A/B of a knob)
TAG=$1; shift
mkdir -p gpurun_out
for kv in "$@"; do
  name=$(echo $kv | tr '= ' '__')
  env $kv python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-secondary > gpurun_out/bench_${TAG}_$name.json 2> gpurun_out/bench_${TAG}_$name.err; echo "== $kv exit $?"; tail -2 gpurun_out/bench_${TAG}_$name.err
  python - <<PY
import json
d=json.loads(open("gpurun_out/bench_${TAG}_$name.json").read().strip().splitlines()[-1])
print("$kv", round(d["ms_per_step"],3), "ms/step", d["clocks"], "tke", d["status"]["tke"], "div", d["status"]["div"])
print("   ", " | ".join(f"{k.replace('_kernel','')} {v['ms_per_launch']}x{v['launches']}" for k,v in list(d["kernels"].items())[:12]))
PY
done
