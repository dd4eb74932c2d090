#!/bin/bash
# round-2 run 1: peaks + round-1 kernels at the BASELINE sizes
bash tools/gpu.sh peaks
export BENCH_TIMEOUT=600
bash tools/gpu.sh line r2a_lampshade_1024 --workload lampshade --steps 2 --warmup 1 --cpu-seconds 5
bash tools/gpu.sh line r2a_cornell_16 --workload cornell --steps 5 --warmup 3 --cpu-seconds 5
bash tools/gpu.sh line r2a_cornell_256 --workload cornell --spp 256 --steps 3 --warmup 3 --cpu-seconds 0
RPTX_PATH_MODE=fused bash tools/gpu.sh line r2a_cornell_256_fused --workload cornell --spp 256 --steps 2 --warmup 1 --cpu-seconds 0
RPTX_PATH_MODE=fused bash tools/gpu.sh line r2a_lampshade_64_fused --workload lampshade --spp 64 --steps 2 --warmup 1 --cpu-seconds 0
bash tools/gpu.sh line r2a_fractal_mb5 --workload fractal_teapots_mb5 --steps 2 --warmup 1 --cpu-seconds 5
for k in photonphoton beambeam beamphoton; do
  bash tools/gpu.sh line r2a_${k}_1e8 --workload $k --photons 1e8 --steps 1 --warmup 1 --cpu-seconds 5
  nvidia-smi --query-gpu=memory.used --format=csv,noheader
done
