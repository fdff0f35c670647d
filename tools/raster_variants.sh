#!/bin/bash
# Forward rasteriser compiled for 4 instead of 5 resident CTAs per SM (54 instead of 48 registers), as a complete
# library next to the product one.   build here: tools/raster_variants.sh build    GPU box: tools/raster_variants.sh run
cd "$(dirname "$0")/.."
V=clm-gs_b200/build_variants
if [ "$1" = build ]; then
  mkdir -p $V
  python clm-gs_b200/build.py > /dev/null || exit 1
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden \
    --expt-relaxed-constexpr -use_fast_math -DCLMGS_RASTER_FWD_MIN_BLOCKS=4 -x cu -c clm-gs_b200/csrc/raster.cu -o $V/raster_fwd4.o || exit 1
  objs=$(ls clm-gs_b200/build/*.o | grep -v "/raster.o")
  nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $V/libclmgs_fwd4.so $objs $V/raster_fwd4.o -lpthread || exit 1
  ls -la $V/libclmgs_fwd4.so
else
  for cfg in c2_bicycle_6m c4_bigcity_102m c3_rubble4k_28m; do
    echo "== $cfg product (5 CTAs/SM)"; python tools/raster_bench.py $cfg
  done
  for cfg in c2_bicycle_6m c4_bigcity_102m; do
    echo "== $cfg forward compiled for 4 CTAs/SM"; CLMGS_LIB_PATH=$PWD/$V/libclmgs_fwd4.so python tools/raster_bench.py $cfg
    echo "== $cfg per-tile shared-memory accumulation in the backward"; CLMGS_RASTER_BWD_TILE_ACC=1 python tools/raster_bench.py $cfg
  done
fi
