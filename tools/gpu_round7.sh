#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
grep -v "^E  *+" gpurun_out/pytest.log | tail -15
for w in cornell lampshade; do
python bench.py --steps 2 --warmup 3 --cpu-seconds 0 --workload $w 2>gpurun_out/b_$w.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['config']['workload'][:30], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), {k:round(v,1) for k,v in d['kernel_ms_per_step'].items()})"
done
