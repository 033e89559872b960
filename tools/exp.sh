# usage: tools/exp.sh TAG "EXTRA flags"   (run on the GPU box: rebuilds the library there with the flags, benches)
TAG=$1; EXTRA=$2
touch microhh_b200/csrc/*.cu
make -s -j16 -C microhh_b200/csrc EXTRA="$EXTRA" >/dev/null 2>&1 || { echo BUILD FAILED; exit 1; }
echo "== $TAG: $EXTRA"
tools/gpu_quick.sh $TAG | head -6
