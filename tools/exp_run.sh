#!/bin/bash
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
tools/gpu_quick.sh lazy 2>/dev/null | head -12
