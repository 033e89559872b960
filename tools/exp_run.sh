#!/bin/bash
python -m pytest tests/test_gpu_ops.py -m gpu -x -q 2>&1 | tail -3
