#!/bin/bash
# Build variants of csrc/radix.cu (pairs per thread x resident CTAs per SM) as stand-alone libraries and time each
# with tools/sort_bench.py.   build here:  tools/radix_sweep.sh build      run on the GPU box:  tools/radix_sweep.sh run
cd "$(dirname "$0")/.."
V=clm-gs_b200/build_variants
ARCH="-gencode arch=compute_100a,code=sm_100a"
FLAGS="-O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr"
VARIANTS="16:3 8:5 8:6 12:4 24:2 32:1"
if [ "$1" = build ]; then
  mkdir -p $V
  for v in $VARIANTS; do
    it=${v%%:*}; bp=${v##*:}
    nvcc $ARCH $FLAGS -DCLMGS_RADIX_ITEMS=$it -DCLMGS_RADIX_BLOCKS_PER_SM=$bp -Xptxas -v -shared -o $V/libradix_${it}_${bp}.so \
      clm-gs_b200/csrc/radix.cu clm-gs_b200/csrc/api.cu 2> $V/ptxas_${it}_${bp}.log || { tail -5 $V/ptxas_${it}_${bp}.log; exit 1; }
    grep -A1 "radix_scatter_kernelILi0" $V/ptxas_${it}_${bp}.log | grep Used | head -1 | sed "s/^/items=$it blocks=$bp: /"
  done
else
  for v in $VARIANTS; do
    it=${v%%:*}; bp=${v##*:}
    echo "== items=$it blocks_per_sm=$bp"
    CLMGS_LIB_PATH=$PWD/$V/libradix_${it}_${bp}.so python tools/sort_bench.py 46000000 13
    CLMGS_LIB_PATH=$PWD/$V/libradix_${it}_${bp}.so python tools/sort_bench.py 25000000 14
    CLMGS_LIB_PATH=$PWD/$V/libradix_${it}_${bp}.so python tools/sort_bench.py 5100000 32
  done
fi
