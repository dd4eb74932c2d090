#!/bin/bash
# multi-GPU with the sharded photon-map build: parity (N ranks == 1 rank, maps bit for bit), then photon scaling at 1e8 with the
# per-phase build trace, sharded against replicated
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multi.log
tail -12 gpurun_out/pytest_multi.log
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
export RPTX_TRACE_BUILD=1
for k in photonphoton beamphoton beambeam; do
  run r2k_${k}_n1 1 --workload $k --photons 1e8 --steps 2 --warmup 1 --cpu-seconds 0
  run r2k_${k}_n$N $N --workload $k --photons 1e8 --steps 2 --warmup 1 --cpu-seconds 0
  grep "rptx\]" gpurun_out/r2k_${k}_n$N.err | tail -14
  RPTX_REPLICATED_MAP_BUILD=1 run r2k_${k}_n${N}_replicated $N --workload $k --photons 1e8 --steps 2 --warmup 1 --cpu-seconds 0
done
unset RPTX_TRACE_BUILD
run r2k_lampshade_n$N $N --workload lampshade --spp 256 --steps 3 --warmup 2 --cpu-seconds 0 --no-extra
