#!/bin/bash
# N-GPU photon bench lines (emission sharded by photon index, rows all-gathered over NCCL, replicated map build, sharded camera pass)
mkdir -p gpurun_out
N=${N:-8}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N "$@"; }
for k in ${KINDS:-photonphoton beambeam}; do
  timeout 900 bash -c "$(declare -f run); N=$N run --steps 2 --warmup 1 --cpu-seconds 0 --workload $k --photons ${PHOTONS:-1e7}" > gpurun_out/s${N}_$k.json 2> gpurun_out/s${N}_$k.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/s${N}_$k.json").read().strip().splitlines()[-1])
    print("N=$N $k", d['metric'], "value=%.2f"%d['value'], "e2e=%.2f"%d['e2e']['value'], "ms/step=%.1f"%d['ms_per_step'], d.get('passes_ms'), d['config'].get('photons'))
except Exception as e:
    print("$k failed", e); print(open("gpurun_out/s${N}_$k.err").read()[-1500:])
PY
done
