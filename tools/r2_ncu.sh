#!/bin/bash
# ncu --set full capture of one workload's kernels: tools/r2_ncu.sh <tag> <kernel regex> <launch-skip> <count> <bench args...>
tag=$1; kre=$2; skip=$3; cnt=$4; shift 4
mkdir -p gpurun_out
CMD="python bench.py --no-extra --cpu-seconds 0 $*"
$CMD > gpurun_out/${tag}_plain.log 2>&1 || { echo plain run failed; tail -5 gpurun_out/${tag}_plain.log; exit 1; }
tail -c 400 gpurun_out/${tag}_plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"$kre" -s $skip -c $cnt -f -o gpurun_out/$tag $CMD > gpurun_out/${tag}_ncu.log 2>&1
tail -3 gpurun_out/${tag}_ncu.log; ls -la gpurun_out/$tag.ncu-rep
