#!/bin/bash
# lagged wave loop: parity suite, issue probes (per-vertex instruction counts), then the default bench line
bash tools/gpu.sh tests
python tools/issue_probe.py lampshade --width 960 --height 540 --spp 32 > gpurun_out/probe.log 2>&1
python tools/issue_probe.py cornell --spp 16 >> gpurun_out/probe.log 2>&1
python tools/issue_probe.py fractal_teapots --spp 16 >> gpurun_out/probe.log 2>&1
python tools/issue_probe.py fractal_teapots_mb5 --spp 4 >> gpurun_out/probe.log 2>&1
RPTX_PATH_MODE=persist python tools/issue_probe.py lampshade --width 960 --height 540 --spp 32 >> gpurun_out/probe.log 2>&1
RPTX_PATH_MODE=persist python tools/issue_probe.py cornell --spp 16 >> gpurun_out/probe.log 2>&1
tail -8 gpurun_out/probe.log | cut -c1-300
cp gpurun_out/r2_issue.json profiles/r2_issue.json
BENCH_TIMEOUT=900 bash tools/gpu.sh line r2d_default --steps 5 --warmup 3
tail -c 1500 gpurun_out/r2d_default.err
