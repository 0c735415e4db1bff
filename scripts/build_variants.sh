#!/bin/bash
# usage: build_variants.sh name="-DFLAG ..." ...  -> point-cloud_b200/variants/libpcb_<name>.so (tuning experiments;
# the GPU-side script copies one over libpcb.so per run)
set -e
cd "$(dirname "$0")/../point-cloud_b200/csrc"
mkdir -p ../variants
[ -f hostio.o ] && [ -f api.o ] || make -s
for spec in "$@"; do
  name=${spec%%=*}; flags=${spec#*=}
  nvcc -gencode arch=compute_100a,code=sm_100a $flags -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC,-Wall,-Wno-unused-function -c builder.cu -o ../variants/builder_$name.o &
done
wait
for spec in "$@"; do
  name=${spec%%=*}
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../variants/libpcb_$name.so ../variants/builder_$name.o hostio.o api.o -cudart shared -Xlinker -rpath,/usr/local/cuda/lib64 -lpthread
  rm ../variants/builder_$name.o
done
ls -la ../variants
