#!/bin/bash
# parity tests + short bench lines (no ncu)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -15 gpurun_out/pytest.log
for w in fractal_teapots cornell lampshade; do
  spp=64; [ $w = cornell ] && spp=256
  timeout 300 python bench.py --steps 2 --warmup 3 --workload $w --spp $spp --cpu-seconds 0 > gpurun_out/q_$w.json 2> gpurun_out/q_$w.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/q_$w.json").read().strip().splitlines()[-1])
    print("$w", "value=%.1f"%d['value'], "e2e=%.1f"%d['e2e']['value'], {k:round(v,1) for k,v in d['kernel_ms_per_step'].items()})
except Exception as e:
    print("$w failed", e); print(open("gpurun_out/q_$w.err").read()[-2000:])
PY
done
