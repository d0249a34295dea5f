#!/bin/bash
# builds libsphb200 variants with different knobs into scratch/variants/<name>.so
set -e
cd "$(dirname "$0")/../sphinxsys-learning_b200/csrc"
mkdir -p ../../scratch/variants
FLAGS="-O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC --expt-relaxed-constexpr"
build() { # name, extra flags
  name=$1; shift
  nvcc $FLAGS "$@" -c fluid.cu -o /tmp/t/fluid_$name.o
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../../scratch/variants/$name.so /tmp/t/fluid_$name.o context.o cell_list.o relation.o solid.o slab.o -lcudart
}
for spec in "$@"; do
  IFS=: read name f64 f32 un <<< "$spec"
  build $name -DSPHB_MINB_F64=$f64 -DSPHB_MINB_F32=$f32 -DSPHB_UNROLL=$un &
done
wait
ls -la ../../scratch/variants
