#!/bin/bash
# block-synchronous persistent kernel: parity, then speed against the wavefront (512x1 and 256x2 blocks per SM)
bash tools/gpu.sh tests tests/test_gpu_render.py
export BENCH_TIMEOUT=300
for lib in "" _q2; do
  [ -n "$lib" ] && export RPTX_LIBRARY=$PWD/rpt_b200/librptx$lib.so
  RPTX_PATH_MODE=persist bash tools/gpu.sh line r2c_lampshade_128_persist$lib --workload lampshade --spp 128 --steps 2 --warmup 1 --cpu-seconds 0 --no-extra
  RPTX_PATH_MODE=persist bash tools/gpu.sh line r2c_cornell_256_persist$lib --workload cornell --spp 256 --steps 2 --warmup 1 --cpu-seconds 0
done
unset RPTX_LIBRARY
RPTX_PATH_MODE=persist bash tools/gpu.sh line r2c_cornell_16_persist --workload cornell --steps 5 --warmup 3 --cpu-seconds 0
