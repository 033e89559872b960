#!/bin/bash
# usage: tools/gpu_a2a.sh TAG N "modes"  -> 2-rank slab parity test per mode, then the bench on N GPUs per all-to-all mode of the peer path
TAG=$1; N=$2; MODES=${3:-"fused fused_pipe ce"}
mkdir -p gpurun_out
for mode in $MODES; do
  echo "######## $mode"
  [ -z "$SKIP_PARITY" ] && MHH_SLAB_A2A=$mode SLAB_DIMS=32,512,8 MHH_SLAB_PATH=peer timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tests/slab_parity.py 2>&1 | grep -E "SLAB PARITY|Error|error" | head -5
  MHH_SLAB_A2A=$mode MHH_SLAB_PATH=peer timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 $BENCH_EXTRA > gpurun_out/bench_${TAG}_n${N}_$mode.json 2> gpurun_out/bench_${TAG}_n${N}_$mode.err; echo "== $mode exit $?"; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/bench_${TAG}_n${N}_$mode.err | tail -4
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${TAG}_n${N}_$mode.json").read().strip().splitlines()[-1])
    print(d.get("slab_exchange_timing")); print("$mode", round(d["ms_per_step"],3), "ms/step", d["clocks"], "parity", d["slab_parity"]["max_abs_diff"], d["slab_parity"]["ok"], "tke", d["status"]["tke"], "div", d["status"]["div"])
    print("   ", " | ".join(f"{k.replace('_kernel','')} {v['ms_per_launch']}x{v['launches']}" for k,v in list(d["kernels"].items())[:14]))
except Exception as e: print("no line", e)
PY
done
