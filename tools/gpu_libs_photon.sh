#!/bin/bash
# photon bench lines of several builds: LIBS="a.so b.so" KINDS="photonphoton beambeam" bash tools/gpu_libs_photon.sh
mkdir -p gpurun_out
for k in ${KINDS:-photonphoton}; do for l in $LIBS; do
RPTX_LIBRARY=$PWD/rpt_b200/$l timeout 600 python bench.py --steps 2 --warmup 1 --workload $k --photons ${PHOTONS:-1e7} --cpu-seconds 0 2>gpurun_out/p_$k.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$l', '$k', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), round(d['ms_per_step'],1), {k:round(v,1) for k,v in d.get('passes_ms').items()})"
done; done
