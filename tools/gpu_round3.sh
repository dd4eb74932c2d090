#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
grep -v "^E  *+" gpurun_out/pytest.log | tail -30
for mode in ${MODES:-wavefront fused}; do for w in ${WORKLOADS:-fractal_teapots cornell lampshade teapot}; do
  spp=64; [ $w = cornell ] && spp=256
  RPTX_PATH_MODE=$mode timeout 300 python bench.py --steps 2 --warmup 3 --workload $w --spp $spp --cpu-seconds 0 > gpurun_out/q_${mode}_$w.json 2> gpurun_out/q_${mode}_$w.err
  python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/q_${mode}_$w.json").read().strip().splitlines()[-1])
    print("$mode $w", "value=%.1f"%d['value'], "e2e=%.1f"%d['e2e']['value'], {k:round(v,1) for k,v in d['kernel_ms_per_step'].items()})
except Exception as e:
    print("$mode $w failed", e); print(open("gpurun_out/q_${mode}_$w.err").read()[-1500:])
PY
done; done
python tools/gather_prof.py photonphoton 1e6 20 2>&1 | cut -c1-60
python tools/gather_prof.py beamphoton 1e6 20 2>&1 | cut -c1-60
python tools/gather_prof.py beambeam 1e7 20 2>&1 | cut -c1-60
