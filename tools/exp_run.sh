#!/bin/bash
python -m pytest tests/test_gpu_ops.py tests/test_gpu_les_substep.py -m gpu -x -q 2>&1 | tail -2
tools/exp1.sh final diff_smag2.cu "" 8
tools/exp_dns.sh base ""
tools/exp_dns.sh u2 "-DMHH_ADVEC4_UNROLL_N=2 -DMHH_ADVEC2I5_UNROLL_N=2 -DMHH_PRES_IN_UNROLL_N=2 -DMHH_DIFF2_UNROLL_N=2"
tools/exp_dns.sh u4 "-DMHH_ADVEC4_UNROLL_N=4 -DMHH_ADVEC2I5_UNROLL_N=4 -DMHH_PRES_IN_UNROLL_N=4 -DMHH_DIFF2_UNROLL_N=4"
tools/exp_dns.sh u3 "-DMHH_ADVEC4_UNROLL_N=3 -DMHH_ADVEC2I5_UNROLL_N=3 -DMHH_PRES_IN_UNROLL_N=8 -DMHH_DIFF2_UNROLL_N=8"
