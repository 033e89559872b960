#!/bin/bash
python -m pytest tests/test_gpu_ops.py tests/test_gpu_les_substep.py tests/test_gpu_multistep.py -m gpu -x -q 2>&1 | tail -3
for V in "" "-DMHH_UVW_ROLES_UNROLL_N=1" "-DMHH_UVW_ROLES_UNROLL_N=2 -DMHH_UVW_ROLES_MINB=1" "-DMHH_UVW_ROLES=0"; do
  tools/exp1.sh roles diff_smag2.cu "$V" 5 | grep "==\|uvw\|^3\|^4"
done
