#!/bin/bash
# usage (on the GPU box): tools/gpu_round.sh TAG [full]  -> gpu tests, bench, dns4 bench, ncu launch list (+ --set full capture of one substep)
TAG=$1; FULL=$2
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_$TAG.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/pytest_$TAG.log
python bench.py --steps 5 --warmup 3 > gpurun_out/bench_512_$TAG.json 2> gpurun_out/bench_512_$TAG.err; echo "bench exit $?"
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_512_$TAG.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["value"], d["e2e"], d["roofline"], d["cpu_baseline"], d["clocks"])
for k,v in list(d["kernels"].items())[:16]: print(k, v["ms_per_launch"], v["gbs"])
PY
python tools/bench_dns4.py > gpurun_out/dns4_$TAG.txt 2>&1; tail -30 gpurun_out/dns4_$TAG.txt
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_$TAG.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launch_$TAG.log 2>&1; echo "ncu launches exit $?"
if [ -n "$FULL" ]; then
  ncu --set full --clock-control none --import-source on -k regex:"evisc_kernel|tile_kernel|pres2_in|fft_pow2|tdma|out_rk|::rk_kernel" -s 10 -c 10 -f -o gpurun_out/prof_${TAG}_step python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_full_$TAG.log 2>&1; echo "ncu full exit $?"
fi
