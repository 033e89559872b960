#!/bin/bash
# usage (GPU box): tools/exp_dns.sh TAG "EXTRA flags" -> rebuild advec_4/advec_2i5/pres/diff_2 with the flags, tools/bench_dns4.py summary
TAG=$1; EXTRA=$2
touch microhh_b200/csrc/advec_4.cu microhh_b200/csrc/advec_2i5.cu microhh_b200/csrc/pres_2.cu microhh_b200/csrc/pres_4.cu microhh_b200/csrc/diff_2.cu
make -s -j16 -C microhh_b200/csrc EXTRA="$EXTRA" >/dev/null 2>&1 || { echo BUILD FAILED; exit 1; }
echo "== $TAG: $EXTRA"
python tools/bench_dns4.py 2>/dev/null | grep "^==\|advec\|pres._in\|diff2" | cut -c1-60
