#!/bin/bash
N=$1
python -m pytest tests/test_gpu_slab.py -m gpu -x -q 2>&1 | tail -3
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n${N}_r01j.json 2> gpurun_out/bench_n${N}_r01j.err; echo "bench n$N exit $?"
python - <<PY
import json
d=json.loads(open("gpurun_out/bench_n${N}_r01j.json").read().strip().splitlines()[-1])
print(d["ms_per_step"], d["value"], d["config"]["workload"], d["clocks"])
PY
