#!/bin/bash
# 8 GPUs: parity (4 ranks), strong scaling of the headline and cornell, photon workloads with the sharded and the replicated map build
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multi.log
tail -5 gpurun_out/pytest_multi.log
run() { # tag, nproc, bench args
  local tag=$1 n=$2; shift 2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n "$@" > gpurun_out/$tag.json 2> gpurun_out/$tag.err
  python - "$tag" <<'PY'
import json, sys
tag = sys.argv[1]
try:
    d = json.loads([l for l in open(f"gpurun_out/{tag}.json").read().strip().splitlines() if l.startswith("{")][-1])
    k = d.get("kernel_ms_per_step") or d.get("passes_ms") or {}
    print(tag, "n=%d" % d["n_gpus"], d["scaling"], "value=%.2f" % d["value"], "e2e=%.2f" % d["e2e"]["value"], "ms/step=%.1f" % d["ms_per_step"], {a: round(b, 1) for a, b in k.items()})
except Exception as e:
    print(tag, "FAILED", e); print(open(f"gpurun_out/{tag}.err").read()[-1500:])
PY
}
run r2m_lampshade_n$N $N --steps 3 --warmup 3 --cpu-seconds 0
run r2m_lampshade_n4 4 --steps 3 --warmup 3 --cpu-seconds 0
run r2m_cornell_n$N $N --workload cornell --steps 10 --warmup 3 --cpu-seconds 0
run r2m_fractal_mb5_n$N $N --workload fractal_teapots_mb5 --steps 3 --warmup 2 --cpu-seconds 0
export RPTX_TRACE_BUILD=1
for k in photonphoton beamphoton beambeam; do
  run r2m_${k}_n$N $N --workload $k --photons 1e8 --steps 2 --warmup 1 --cpu-seconds 0
  grep "rptx\]" gpurun_out/r2m_${k}_n$N.err | grep -v "emit:" | tail -12
  RPTX_REPLICATED_MAP_BUILD=1 run r2m_${k}_n${N}_replicated $N --workload $k --photons 1e8 --steps 2 --warmup 1 --cpu-seconds 0
done
