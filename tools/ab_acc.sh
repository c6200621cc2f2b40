#!/bin/bash
# A/B of the KMeans filter kernel variants on one GPU: product library vs build_variants/<name>
cd "$(dirname "$0")/.."
for shape in "--n 20000000 --d 128 --k 1024" "--n 40000000 --d 64 --k 256"; do
  for lib in "" $@; do
    if [ -z "$lib" ]; then echo "== product $shape"; python tools/km_bench.py $shape --algo 5 --reps 5 --lloyd 2
    else echo "== $lib $shape"; B2F_LIB_PATH=build_variants/$lib/libb200fit.so python tools/km_bench.py $shape --algo 5 --reps 5 --lloyd 2; fi
  done
done
