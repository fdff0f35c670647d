#!/bin/bash
# GPU box: parity tests that touch the cull (bit-exact masks / ids against the oracle and the reference's kernel),
# then the cull alone (count + scan + emit) on prefixes of the three bench scenes.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests/test_gpu_forward.py tests/test_gpu_vs_reference_cuda.py tests/test_gpu_fullsize_properties.py \
  tests/test_gpu_engine.py tests/test_gpu_api_mirrors.py -m gpu -q > gpurun_out/pytest_cull.log 2>&1; tail -2 gpurun_out/pytest_cull.log
{ python tools/cull_bench.py c4_bigcity_102m 20000000 64
  python tools/cull_bench.py c3_rubble4k_28m 20000000 16
  python tools/cull_bench.py c2_bicycle_6m 6000000 16
  python tools/cull_bench.py c2_bicycle_6m 6000000 4; } 2>&1 | grep -v Warning | tee gpurun_out/cull_ab.txt
