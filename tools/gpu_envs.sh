#!/bin/bash
# bench lines of one build under several values of one env knob: VAR=NAME VALUES="a b c" WORKLOADS="cornell:256" bash tools/gpu_envs.sh
mkdir -p gpurun_out
run() {
python bench.py --steps 2 --warmup 3 --cpu-seconds 0 --workload $1 --spp $2 2>gpurun_out/b_$1.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$3', d['config']['workload'][:10], round(d['value'],1), {k:round(v,1) for k,v in d['kernel_ms_per_step'].items() if k!='wall'})"
}
for w in ${WORKLOADS:-cornell:256 lampshade:64}; do
  for v in $VALUES; do env $VAR=$v bash -c "$(declare -f run); run ${w%%:*} ${w##*:} $VAR=$v"; done
done
