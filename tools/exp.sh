set -e
for cfg in "4 32 0" "4 32 1" "4 32 2" "4 32 4" "4 64 2" "5 32 2"; do
  set -- $cfg
  make -s -C microhh_b200/csrc clean >/dev/null; make -s -j16 -C microhh_b200/csrc EXTRA="-DMHH_MINB=$1 -DMHH_KCHUNK=$2 -DMHH_PF=$3" >/dev/null 2>&1
  python bench.py --steps 4 --warmup 2 --no-cpu-baseline > gpurun_out/exp_$1_$2_$3.json 2>/dev/null
  python - <<PY
import json; d=json.load(open("gpurun_out/exp_$1_$2_$3.json"))
print("minb=$1 kchunk=$2 pf=$3 ms/step", round(d["ms_per_step"],2), {k[:14]: round(v["ms_per_launch"],2) for k,v in d["kernels"].items() if v["share"]>0.03})
PY
done
