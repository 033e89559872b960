#!/bin/bash
# GPU-box experiment: parity of the current build, then bench variants of the momentum kernel
python -m pytest tests/test_gpu_ops.py tests/test_gpu_les_substep.py tests/test_gpu_multistep.py -m gpu -x -q 2>&1 | tail -3
tools/gpu_quick.sh base | head -8
tools/exp.sh unroll3 -DMHH_UVW_UNROLL_N=3
tools/exp.sh unroll2 -DMHH_UVW_UNROLL_N=2
