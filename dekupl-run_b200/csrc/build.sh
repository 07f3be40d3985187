#!/bin/bash
# Builds libdekupl_b200.so (sm_100a only) in-tree. -fmad=false: the statistical
# kernels must not contract a*b+c into FMA (the reference's x86-64 build does not).
set -e
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-O2 -fmad=false -diag-suppress 550"
OBJS=""
for f in dk_ctx dk_join dk_nf dk_ttest dk_sort dk_bh dk_filter dk_mask dk_format dk_ingest dk_tags dk_api dk_comm; do
  if [ ! -f build/$f.o ] || [ $f.cu -nt build/$f.o ] || [ -n "$(find . -maxdepth 1 -name '*.cuh' -newer build/$f.o)" ] || [ ../../include/dekupl_b200.h -nt build/$f.o ]; then
    mkdir -p build
    $NVCC $FLAGS $EXTRA -c $f.cu -o build/$f.o &
  fi
  OBJS="$OBJS build/$f.o"
done
wait
$NVCC -shared -o ../libdekupl_b200.so $OBJS -cudart static -ldl
echo "built $(cd ..; pwd)/libdekupl_b200.so"
