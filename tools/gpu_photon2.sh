#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x --timeout 600 -k "photon or gather or emit" > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
grep -v "^E  *+" gpurun_out/pytest.log | tail -${TAILN:-6}
timeout 1200 python tools/photon_scale.py ${SIZES:-1e7} 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: print(l[:300]); continue
    print(d['variant'], d['photons'], 'emit ms %.1f (%.1f Mph/s, %.2f Gray/s)'%(d['emit_dev_ms'], d['mphotons_per_s'], d['rays']/d['emit_dev_ms']/1e6), 'build ms %.1f'%d['build_dev_ms'], 'gather ms %.1f (%.2f Msamples/s)'%(d['gather_dev_ms'], d['msamples_per_s']), 'mean %.6g'%d['mean'])
"
