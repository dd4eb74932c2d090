#!/bin/bash
mkdir -p gpurun_out
W=${1:-cornell}; SPP=${2:-32}; SKIP=${3:-8}
CMD="python bench.py --steps 1 --warmup 1 --cpu-seconds 0 --spp $SPP --workload $W"
$CMD > gpurun_out/plainc.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_shadow|k_shade|k_connect' -s $SKIP -c 4 -o gpurun_out/prof_$W $CMD > gpurun_out/ncu_c.log 2>&1
tail -2 gpurun_out/ncu_c.log
