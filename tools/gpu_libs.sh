#!/bin/bash
# bench lines of several builds of the library: LIBS="a.so b.so" WORKLOADS="cornell:256 ..." bash tools/gpu_libs.sh
mkdir -p gpurun_out
run() {
python bench.py --steps 2 --warmup 3 --cpu-seconds 0 --workload $1 --spp $2 2>gpurun_out/b_$1.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$3', d['config']['workload'][:10], round(d['value'],1), {k:round(v,1) for k,v in d['kernel_ms_per_step'].items() if k!='wall'})"
}
for w in ${WORKLOADS:-cornell:256 lampshade:64}; do
  for l in $LIBS; do RPTX_LIBRARY=$PWD/rpt_b200/$l run ${w%%:*} ${w##*:} $l; done
done
