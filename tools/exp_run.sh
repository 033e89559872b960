#!/bin/bash
python -m pytest tests/test_gpu_slab.py tests/test_gpu_fft.py -m gpu -x -q 2>&1 | tail -4
