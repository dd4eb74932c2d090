#!/bin/bash
# A/B of an env switch on the two analytic scenes: bash tools/gpu_ab.sh VAR
mkdir -p gpurun_out
run() {
python bench.py --steps 2 --warmup 3 --cpu-seconds 0 --workload $1 --spp $2 2>gpurun_out/b_$1.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$3', d['config']['workload'][:12], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:round(v,1) for k,v in d['kernel_ms_per_step'].items()}, {k:round(v/1e9,2) for k,v in d['rays_per_s'].items()})"
}
for w in "cornell 256" "lampshade 64"; do
  run $w new
  env $1=1 bash -c "$(declare -f run); run $w old"
done
