set -e
make -s -C microhh_b200/csrc clean >/dev/null; make -s -j16 -C microhh_b200/csrc FMAD=true >/dev/null 2>&1
python -m pytest tests -m gpu -q > gpurun_out/fmad_tests.log 2>&1 || true
tail -15 gpurun_out/fmad_tests.log
tools/gpu_quick.sh fmad
