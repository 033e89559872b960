#!/bin/bash
for V in "-DMHH_ADVEC4_MINB=3" "-DMHH_ADVEC4_MINB=4"; do
touch microhh_b200/csrc/advec_4.cu
make -s -j16 -C microhh_b200/csrc EXTRA="$V" >/dev/null 2>&1 || echo BUILD FAILED
echo "== $V"
python tools/bench_dns4.py 2>/dev/null | grep "^== dns4\|advec4" | cut -c1-70
done
