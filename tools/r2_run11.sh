#!/bin/bash
# ncu --set full of the four wavefront kernels in the steady state of a lampshade batch (tile-sorted k_shade)
bash tools/r2_ncu.sh r2_lamp_wave 'k_extend|k_shade|k_shadow|k_connect' 120 8 --workload lampshade --width 1920 --height 1080 --spp 32 --steps 1 --warmup 1
bash tools/gpu.sh line r2h_corn --workload cornell --steps 10 --warmup 3 --cpu-seconds 0 --no-extra
