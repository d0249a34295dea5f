#!/bin/bash
# builds libsphb200 variants into scratch/variants/<name>.so ; usage: build_variants.sh name:"-DFLAG=.. -DFLAG2=.." ...
set -e
cd "$(dirname "$0")/../sphinxsys-learning_b200/csrc"
mkdir -p ../../scratch/variants /tmp/t
FLAGS="-O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC --expt-relaxed-constexpr"
for spec in "$@"; do
  name="${spec%%:*}"; extra="${spec#*:}"
  (
    mkdir -p /tmp/t/$name
    for f in context cell_list relation fluid solid slab ck peer; do nvcc $FLAGS $extra -c $f.cu -o /tmp/t/$name/$f.o & done; wait
    nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../scratch/variants/$name.so /tmp/t/$name/*.o -lcudart
  ) &
done
wait
ls -la ../../scratch/variants
