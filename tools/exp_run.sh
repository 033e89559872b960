#!/bin/bash
echo "== col kernel"; python tools/bench_fft_chunks.py 2>&1 | tail -6
echo "== staged kernel"; MHH_FFT_NOCOL=1 python tools/bench_fft_chunks.py 2>&1 | tail -6
