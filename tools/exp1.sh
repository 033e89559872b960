#!/bin/bash
# usage (GPU box): tools/exp1.sh TAG FILE.cu "EXTRA flags"  -> rebuild one translation unit with the flags, bench, top kernels
TAG=$1; F=$2; EXTRA=$3
touch microhh_b200/csrc/$F
make -s -j16 -C microhh_b200/csrc EXTRA="$EXTRA" >/dev/null 2>&1 || { echo BUILD FAILED; exit 1; }
echo "== $TAG: $EXTRA"
tools/gpu_quick.sh $TAG 2>/dev/null | head -${4:-5}
