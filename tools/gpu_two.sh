#!/bin/bash
# pytest -m gpu, then the two analytic scenes' bench lines (short)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 600 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
grep -v "^E  *+" gpurun_out/pytest.log | tail -${TAILN:-6}
run() {
python bench.py --steps 2 --warmup 3 --cpu-seconds 0 --workload $1 --spp $2 2>gpurun_out/b_$1.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['config']['workload'][:12], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:round(v,1) for k,v in d['kernel_ms_per_step'].items()}, {k:round(v/1e9,2) for k,v in d['rays_per_s'].items()})"
}
for w in ${WORKLOADS:-cornell:256 lampshade:64}; do run ${w%%:*} ${w##*:}; done
