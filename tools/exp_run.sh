#!/bin/bash
python -m pytest tests/test_gpu_schemes.py tests/test_gpu_ops.py tests/test_gpu_golden.py tests/test_gpu_multistep.py -m gpu -x -q 2>&1 | tail -3
python tools/bench_dns4.py 2>/dev/null | grep "^== dns4\|advec4" | cut -c1-70
