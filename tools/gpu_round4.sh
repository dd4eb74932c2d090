#!/bin/bash
mkdir -p gpurun_out
MODES=wavefront bash tools/gpu_round3.sh
CMD="python bench.py --steps 1 --warmup 1 --cpu-seconds 0 --spp 32 --workload cornell"
RPTX_PATH_MODE=wavefront $CMD > gpurun_out/plainc.log 2>&1 && \
RPTX_PATH_MODE=wavefront ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_shadow|k_shade|k_connect' -s 60 -c 4 -o gpurun_out/prof_cornell $CMD > gpurun_out/ncu_c.log 2>&1
tail -2 gpurun_out/ncu_c.log
