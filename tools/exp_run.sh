#!/bin/bash
python -m pytest tests/test_gpu_ops.py -m gpu -x -q -k "fused_pressure_input" 2>&1 | tail -3
for M in 0 1 2; do echo "== MHH_FUSED_PIN=$M"; MHH_FUSED_PIN=$M tools/gpu_quick.sh pin$M 2>/dev/null | head -12 | grep "^[34]\|pres2_in\|fft_fwd_x"; done
