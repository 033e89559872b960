#!/bin/bash
python -m pytest tests/test_gpu_fft.py tests/test_gpu_slab.py tests/test_gpu_ops.py tests/test_gpu_les_substep.py tests/test_gpu_schemes.py -m gpu -x -q 2>&1 | tail -3
tools/gpu_quick.sh colfft 2>/dev/null | head -14
MHH_FFT_NOCOL=1 tools/gpu_quick.sh nocol 2>/dev/null | head -14 | grep "fft\|^4"
