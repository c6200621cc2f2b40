#!/bin/bash
# Dev tool: build a variant of the whole library with extra -D flags applied to EVERY source file.
#   tools/build_variant_all.sh NAME -DFLAG ...   ->  build_variants/NAME/libb200fit.so  (use with B2F_LIB_PATH)
set -e
ROOT=$(cd "$(dirname "$0")/.." && pwd)
NAME=$1; shift
OUT=$ROOT/build_variants/$NAME
mkdir -p "$OUT"
CS=$ROOT/orange3-spark_b200/csrc
for f in "$CS"/*.cu; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -I"$ROOT/include" -I"$CS" "$@" \
    -c "$f" -o "$OUT/$(basename "${f%.cu}").o" &
done
wait
nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o "$OUT/libb200fit.so" "$OUT"/*.o
echo "$OUT/libb200fit.so"
