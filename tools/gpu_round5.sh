#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
grep -v "^E  *+" gpurun_out/pytest.log | tail -25
python tools/gather_prof.py photonphoton 1e6 20 2>&1 | cut -c1-70
RPTX_GATHER_PER_THREAD=1 python tools/gather_prof.py photonphoton 1e6 20 2>&1 | cut -c1-70
python tools/gather_prof.py photonphoton 1e7 100 2>&1 | cut -c1-70
