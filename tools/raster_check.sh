#!/bin/bash
# GPU box: parity tests that touch the rasteriser (both backward variants), kernel times on one camera of each
# config, and one ncu --set full capture of the raster kernels (raw CSV only; the report stays on the box).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_gpu_vs_reference_cuda.py tests/test_gpu_backward.py tests/test_gpu_fullsize_properties.py \
  tests/test_gpu_api_mirrors.py tests/test_gpu_engine.py -m gpu -q > gpurun_out/pytest_raster2.log 2>&1; tail -2 gpurun_out/pytest_raster2.log
CLMGS_RASTER_BWD_TILE_ACC=1 python -m pytest tests/test_gpu_vs_reference_cuda.py tests/test_gpu_backward.py -m gpu -q -k raster \
  > gpurun_out/pytest_raster_ta1.log 2>&1; tail -2 gpurun_out/pytest_raster_ta1.log
for cfg in c2_bicycle_6m c4_bigcity_102m c3_rubble4k_28m; do
  python tools/raster_bench.py $cfg; CLMGS_RASTER_BWD_TILE_ACC=1 python tools/raster_bench.py $cfg
done > gpurun_out/raster_ab2.jsonl 2>&1
cut -c1-230 gpurun_out/raster_ab2.jsonl
ncu --set full --clock-control none --import-source on -k regex:raster_bwd -c 2 -o /tmp/raster_prof python tools/raster_bench.py c2_bicycle_6m \
  > gpurun_out/ncu_raster.log 2>&1
ncu -i /tmp/raster_prof.ncu-rep --page raw --csv > gpurun_out/raster_prof_raw.csv 2>/dev/null
ncu --set full --clock-control none --import-source on -k regex:raster_fwd -c 2 -o /tmp/raster_prof_f python tools/raster_bench.py c2_bicycle_6m \
  >> gpurun_out/ncu_raster.log 2>&1
ncu -i /tmp/raster_prof_f.ncu-rep --page raw --csv > gpurun_out/raster_fwd_prof_raw.csv 2>/dev/null
ls -la gpurun_out/raster_prof_raw.csv
