#!/bin/bash
# Build a variant of libmtsb200.so with extra -D flags into _exp/<name>.so (kernel experiments;
# select it at run time with MTSB200_LIB=_exp/<name>.so).  Usage: tools/build_variant.sh name "-DFOO=1 -DBAR=2"
set -e
cd "$(dirname "$0")/.."
name=$1; flags=$2
mkdir -p _exp/build_$name
NV="/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-Wall,-Wno-unused-function --expt-relaxed-constexpr $flags"
pids=()
for f in api bvh_build trace query render multi noalloc; do
  $NV -c mitsuba_b200/csrc/$f.cu -o _exp/build_$name/$f.o & pids+=($!)
done
for p in "${pids[@]}"; do wait $p; done
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o _exp/$name.so _exp/build_$name/*.o -lcudart -ldl
rm -rf _exp/build_$name
echo built _exp/$name.so
