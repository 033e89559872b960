#!/bin/bash
tools/exp1.sh k64 diff_smag2.cu "" 6
tools/exp1.sh k128 diff_smag2.cu "-DMHH_TILE_KCHUNK=128" 6
python -m pytest tests/test_gpu_ops.py tests/test_gpu_les_substep.py -m gpu -x -q 2>&1 | tail -2
tools/exp1.sh k256 diff_smag2.cu "-DMHH_TILE_KCHUNK=256" 6
