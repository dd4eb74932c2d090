#!/bin/bash
# PLOC object trees: tree-equality test + intersect parity under PLOC, then fractal / teapot lines per builder
timeout 600 python -m pytest tests/test_gpu_render.py -m gpu -x -q -k "sah_trees" 2>&1 | tail -8
RPTX_BVH_BUILDER=ploc timeout 900 python -m pytest tests/test_gpu_intersect.py tests/test_gpu_quirks.py -m gpu -x -q 2>&1 | tail -6
bash tools/gpu.sh tests
export BENCH_TIMEOUT=300
for b in host ploc; do
  export RPTX_BVH_BUILDER=$b
  bash tools/gpu.sh line r2s_mb5_$b --workload fractal_teapots_mb5 --spp 64 --steps 3 --warmup 2 --cpu-seconds 0 --no-extra
  bash tools/gpu.sh line r2s_mb0_$b --workload fractal_teapots --spp 64 --steps 3 --warmup 2 --cpu-seconds 0 --no-extra
  bash tools/gpu.sh line r2s_teapot_$b --workload teapot --steps 3 --warmup 2 --cpu-seconds 0 --no-extra
done
unset RPTX_BVH_BUILDER
RPTX_LBVH_ONLY=1 bash tools/gpu.sh line r2s_mb5_lbvh --workload fractal_teapots_mb5 --spp 64 --steps 3 --warmup 2 --cpu-seconds 0 --no-extra
