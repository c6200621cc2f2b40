#!/bin/bash
# Dev tool: build a variant of the library with extra -D flags for kmeans_tcb.cu / lr_tc.cu experiments.
#   tools/build_variant.sh NAME FILE.cu -DFLAG ...   ->  build_variants/NAME/libb200fit.so  (use with B2F_LIB_PATH)
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
NAME=$1; FILE=$2; shift 2
OUT=$ROOT/build_variants/$NAME
mkdir -p "$OUT"
CS=$ROOT/orange3-spark_b200/csrc
make -C "$CS" >/dev/null
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -I"$ROOT/include" -I"$CS" "$@" \
  -c "$CS/$FILE" -o "$OUT/variant.o"
OBJS=$(ls "$CS"/build/*.o | grep -v "/${FILE%.cu}.o")
nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o "$OUT/libb200fit.so" $OBJS "$OUT/variant.o"
echo "$OUT/libb200fit.so"
