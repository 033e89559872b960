set -e
make -s -C microhh_b200/csrc clean >/dev/null; make -s -j16 -C microhh_b200/csrc FMAD=true >/dev/null 2>&1
python -m pytest tests/test_gpu_les_substep.py tests/test_gpu_schemes.py tests/test_gpu_multistep.py -m gpu -q > gpurun_out/fmad_tests.log 2>&1 || true
tail -15 gpurun_out/fmad_tests.log
python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_fmad.json 2>/dev/null
