#!/bin/bash
# monomial-surface parity tests + a fractal bench line (regression check of k_extend)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "monomial" > gpurun_out/pytest_mono.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_mono.log
tail -25 gpurun_out/pytest_mono.log
timeout 300 python bench.py --steps 2 --warmup 3 --workload fractal_teapots --spp 64 --cpu-seconds 0 > gpurun_out/q_fr.json 2> gpurun_out/q_fr.err
python - <<PY
import json
d=json.loads(open("gpurun_out/q_fr.json").read().strip().splitlines()[-1])
print("fractal value=%.1f"%d['value'], {k:round(v,1) for k,v in d['kernel_ms_per_step'].items()})
PY
