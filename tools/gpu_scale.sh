#!/bin/bash
# N-GPU bench lines of the three path-tracing workloads: N=8 bash tools/gpu_scale.sh
mkdir -p gpurun_out
N=${N:-8}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@"; }
for w in ${WORKLOADS:-fractal_teapots:256 cornell:256 lampshade:64}; do
  name=${w%%:*}; spp=${w##*:}
  timeout 900 bash -c "$(declare -f run); N=$N run --steps 2 --warmup 3 --cpu-seconds 0 --workload $name --spp $spp" > gpurun_out/s${N}_$name.json 2> gpurun_out/s${N}_$name.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/s${N}_$name.json").read().strip().splitlines()[-1])
    print("N=$N $name", "value=%.1f"%d['value'], "e2e=%.1f"%d['e2e']['value'], "ms/step=%.1f"%d['ms_per_step'], {k:round(v,1) for k,v in d['kernel_ms_per_step'].items()})
except Exception as e:
    print("$name failed", e); print(open("gpurun_out/s${N}_$name.err").read()[-1500:])
PY
done
