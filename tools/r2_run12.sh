#!/bin/bash
# steady-state ncu capture (waves 5-6 of a 128-spp lampshade batch), sanitizer runs, then the default bench line
bash tools/r2_ncu.sh r2_lamp_wave 'k_extend|k_shade|k_shadow|k_connect' 20 8 --workload lampshade --width 1920 --height 1080 --spp 128 --steps 1 --warmup 1
bash tools/sanitize.sh
BENCH_TIMEOUT=900 bash tools/gpu.sh line r2i_default --steps 3 --warmup 3
tail -c 600 gpurun_out/r2i_default.err
