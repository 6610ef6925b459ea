#!/bin/bash
# Builds a variant of libgsr_b200.so with extra nvcc defines into dynamic3dgaussians_b200/variants/<name>.so
# (A/B runs: GSR_LIB_PATH=dynamic3dgaussians_b200/variants/<name>.so python bench.py).  usage: tests/build_variant.sh name -DX=1 ...
set -e
cd "$(dirname "$0")/../dynamic3dgaussians_b200"
name=$1; shift
mkdir -p variants build_$name
for f in abi preprocess radix_sort binning blend_forward blend_backward blend_features preprocess_backward peer_allreduce knn dyn_losses image_loss adam train_glue; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr "$@" -c csrc/$f.cu -o build_$name/$f.o &
done
wait
nvcc -shared -o variants/$name.so build_$name/*.o -gencode arch=compute_100a,code=sm_100a -lcudart
rm -rf build_$name
echo variants/$name.so
