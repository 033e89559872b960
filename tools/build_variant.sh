#!/bin/bash
# usage: tools/build_variant.sh NAME FILE.cu "FLAGS"  -> microhh_b200/variants/libmhh_b200_NAME.so (FILE.cu rebuilt with FLAGS, the rest from build/)
set -e
NAME=$1; FILE=$2; FLAGS=$3
cd "$(dirname "$0")/../microhh_b200/csrc"
mkdir -p build_var/$NAME ../variants
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC -Xcompiler -fvisibility=hidden $FLAGS -c $FILE -o build_var/$NAME/${FILE%.cu}.o
OBJS=$(ls build/*.o | grep -v "/${FILE%.cu}.o")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/libmhh_b200_$NAME.so $OBJS build_var/$NAME/${FILE%.cu}.o -lcudart
echo built $NAME
