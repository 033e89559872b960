#!/bin/bash
# usage: tools/gpu_variants.sh TAG  -> full GPU test suite, then the 512^3 bench under several build / runtime variants
TAG=$1
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --durations=8 > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest exit $?"; tail -25 gpurun_out/pytest_$TAG.log
run() {  # name, env...
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
run notma MHH_TMA=0
run nofuseforce MHH_FUSE_FORCE=0
[ -f microhh_b200/variants/libmhh_b200_fmad.so ] && run fmad MHH_B200_LIB=$PWD/microhh_b200/variants/libmhh_b200_fmad.so
