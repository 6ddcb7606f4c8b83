#!/bin/bash
# builds tuning variants of libsvo_gpu.so into svo_b200/lib/variants/<name>.so:  scripts/build_variants.sh name "flags" [name "flags" ...]
set -e
mkdir -p svo_b200/lib/variants
while [ $# -ge 2 ]; do
  name=$1; flags=$2; shift 2
  rm -rf /tmp/svo_variant_obj; mkdir -p /tmp/svo_variant_obj
  for f in pyramid fast bucket klt klt_generic filter capi; do
    /usr/local/cuda/bin/nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Xcompiler -fPIC,-fvisibility=hidden -Iinclude $flags -Xptxas -v -c svo_b200/csrc/$f.cu -o /tmp/svo_variant_obj/$f.o 2>&1 | grep -A2 "klt_run_kernelILi21" | grep Used | sed "s/^/$name: /" &
  done
  wait
  /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -cudart static -o svo_b200/lib/variants/$name.so /tmp/svo_variant_obj/*.o
done
ls -la svo_b200/lib/variants
