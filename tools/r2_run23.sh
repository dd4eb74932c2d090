#!/bin/bash
# PLOC search radius sweep on fractal_teapots mb5 (extend time = tree quality), then the photon gather ncu captures
export BENCH_TIMEOUT=300 RPTX_BVH_BUILDER=ploc
for r in 8 16 32 64 128; do
  RPTX_PLOC_RADIUS=$r bash tools/gpu.sh line r2t_mb5_r$r --workload fractal_teapots_mb5 --spp 64 --steps 3 --warmup 2 --cpu-seconds 0 --no-extra
done
unset RPTX_BVH_BUILDER
bash tools/r2_run22.sh
