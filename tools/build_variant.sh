#!/bin/bash
# build a variant of the library with extra -D flags for ONE source file (A/B runs on the GPU box via DK_LIB)
# usage: tools/build_variant.sh <name> <file-without-.cu> "<flags>"   -> build/var/lib_<name>.so
set -e
NAME=$1; F=$2; FLAGS_X=$3
cd "$(dirname "$0")/../dekupl-run_b200/csrc"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-O2 -fmad=false -diag-suppress 550"
mkdir -p ../../build/var
$NVCC $FLAGS $FLAGS_X -c $F.cu -o ../../build/var/${F}_$NAME.o 2>&1 | grep -v deprecated || true
OBJS=""
for f in dk_ctx dk_join dk_nf dk_ttest dk_sort dk_bh dk_filter dk_mask dk_format dk_ingest dk_tags dk_api dk_comm; do
  if [ $f = $F ]; then OBJS="$OBJS ../../build/var/${F}_$NAME.o"; else OBJS="$OBJS build/$f.o"; fi
done
$NVCC -shared -o ../../build/var/lib_$NAME.so $OBJS -cudart static -ldl 2>&1 | grep -v deprecated || true
echo built build/var/lib_$NAME.so
