#!/bin/bash
# round-2 run 2: the persistent path kernel against the wavefront, occupancy variants
bash tools/gpu.sh tests tests/test_gpu_render.py
export BENCH_TIMEOUT=300
for lib in "" _p3 _p2 _p5; do
  [ -n "$lib" ] && export RPTX_LIBRARY=$PWD/rpt_b200/librptx$lib.so
  [ -n "$lib" ] && [ ! -f "$RPTX_LIBRARY" ] && continue
  RPTX_PATH_MODE=persist bash tools/gpu.sh line r2b_lampshade_128_persist$lib --workload lampshade --spp 128 --steps 2 --warmup 1 --cpu-seconds 0
  RPTX_PATH_MODE=persist bash tools/gpu.sh line r2b_cornell_256_persist$lib --workload cornell --spp 256 --steps 2 --warmup 1 --cpu-seconds 0
done
unset RPTX_LIBRARY
RPTX_PATH_MODE=persist bash tools/gpu.sh line r2b_cornell_16_persist --workload cornell --steps 5 --warmup 3 --cpu-seconds 0
RPTX_PATH_MODE=persist bash tools/gpu.sh line r2b_fractal_mb5_persist --workload fractal_teapots_mb5 --spp 32 --steps 2 --warmup 1 --cpu-seconds 0
RPTX_PATH_MODE=persist bash tools/gpu.sh line r2b_fractal_mb0_persist --workload fractal_teapots --spp 64 --steps 2 --warmup 1 --cpu-seconds 0
