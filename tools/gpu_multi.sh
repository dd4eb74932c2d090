#!/bin/bash
# 2-GPU check of the NCCL paths (path tracing all-reduce, photon all-gather)
mkdir -p gpurun_out
N=${N:-2}
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N "$@"; }
timeout 600 python bench.py --steps 2 --warmup 3 --cpu-seconds 0 > gpurun_out/m1_fractal.json 2> gpurun_out/m1_fractal.err
timeout 600 bash -c "$(declare -f run); N=$N run --steps 2 --warmup 3 --cpu-seconds 0" > gpurun_out/m${N}_fractal.json 2> gpurun_out/m${N}_fractal.err
timeout 600 python bench.py --steps 2 --warmup 1 --cpu-seconds 0 --workload photonphoton --photons 1e7 > gpurun_out/m1_pp.json 2> gpurun_out/m1_pp.err
timeout 600 bash -c "$(declare -f run); N=$N run --steps 2 --warmup 1 --cpu-seconds 0 --workload photonphoton --photons 1e7" > gpurun_out/m${N}_pp.json 2> gpurun_out/m${N}_pp.err
timeout 600 bash -c "$(declare -f run); N=$N run --steps 2 --warmup 1 --cpu-seconds 0 --workload beambeam --photons 1e7" > gpurun_out/m${N}_bb.json 2> gpurun_out/m${N}_bb.err
timeout 600 bash -c "$(declare -f run); N=$N run --steps 2 --warmup 3 --cpu-seconds 0 --workload cornell --spp 256" > gpurun_out/m${N}_cornell.json 2> gpurun_out/m${N}_cornell.err
for f in m1_fractal m${N}_fractal m1_pp m${N}_pp m${N}_bb m${N}_cornell; do
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/$f.json").read().strip().splitlines()[-1])
    print("$f", d['metric'], "value=%.2f"%d['value'], "e2e=%.2f"%d['e2e']['value'], "ms/step=%.1f"%d['ms_per_step'], d.get('passes_ms') or d.get('kernel_ms_per_step'))
except Exception as e:
    print("$f failed", e); print(open("gpurun_out/$f.err").read()[-1500:])
PY
done
