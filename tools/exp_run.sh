#!/bin/bash
# ncu --set full of the 4th-order DNS and the advec_2i5 LES kernels (one launch each of the first captured substeps)
ncu --set full --clock-control none --import-source on -k regex:"advec4_mom|advec4_diff4|pres4_|advec2i5|diff2_uvw" -s 12 -c 12 -f -o gpurun_out/prof_r01i_dns4 python tools/bench_dns4.py > gpurun_out/ncu_dns4.log 2>&1; echo "ncu exit $?"; tail -3 gpurun_out/ncu_dns4.log
