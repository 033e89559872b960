#!/bin/bash
# usage: tools/gpu_slab.sh TAG N [paths]  -> slab tests, then the bench on N GPUs with the given slab paths (default: peer nccl)
TAG=$1; N=$2; PATHS=${3:-"peer nccl"}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_slab.py tests/test_gpu_les_substep.py tests/test_gpu_parity_large.py -m gpu -q > gpurun_out/pytest_slab_$TAG.log 2>&1; echo "pytest exit $?"; grep -E "slab parity|SLAB PARITY|passed|failed|Error" gpurun_out/pytest_slab_$TAG.log | tail -30
for path in $PATHS; do
  MHH_SLAB_PATH=$path timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/bench_${TAG}_n${N}_$path.json 2> gpurun_out/bench_${TAG}_n${N}_$path.err; echo "== $path exit $?"; grep -v "^\*\|OMP_NUM\|^$" gpurun_out/bench_${TAG}_n${N}_$path.err | tail -4
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/bench_${TAG}_n${N}_$path.json").read().strip().splitlines()[-1])
    print("$path", round(d["ms_per_step"],3), "ms/step", d["clocks"], d.get("slab_parity"), "tke", d["status"]["tke"], "div", d["status"]["div"])
    print("   ", " | ".join(f"{k.replace('_kernel','')} {v['ms_per_launch']}x{v['launches']}" for k,v in list(d["kernels"].items())[:16]))
except Exception as e: print("no line", e)
PY
done
