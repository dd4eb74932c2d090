#!/bin/bash
# k_shade tile sort: parity suite, then lampshade (256 spp, same per-wave sizes as 1024) and cornell at tile rounds 1 / 4 / 8 / auto
bash tools/gpu.sh tests
export BENCH_TIMEOUT=300
for r in 1 2 4 8 0; do
  export RPTX_SHADE_TILE_ROUNDS=$r
  bash tools/gpu.sh line r2g_lamp_r$r --workload lampshade --spp 256 --steps 3 --warmup 3 --cpu-seconds 0 --no-extra
  bash tools/gpu.sh line r2g_corn_r$r --workload cornell --steps 10 --warmup 3 --cpu-seconds 0 --no-extra
done
unset RPTX_SHADE_TILE_ROUNDS
python tools/issue_probe.py lampshade --width 960 --height 540 --spp 32 > gpurun_out/probe.log 2>&1
python tools/issue_probe.py cornell --spp 16 >> gpurun_out/probe.log 2>&1
tail -4 gpurun_out/probe.log | cut -c1-400
