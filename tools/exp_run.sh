#!/bin/bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -2
tools/gpu_quick.sh cyc 2>/dev/null | grep "^[34]\|cyclic"
