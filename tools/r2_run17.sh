#!/bin/bash
# forward NEE accumulation (k_connect touches finished paths only): parity suite, lampshade / cornell lines, issue probe
bash tools/gpu.sh tests
export BENCH_TIMEOUT=300
bash tools/gpu.sh line r2n_lamp --workload lampshade --spp 256 --steps 3 --warmup 3 --cpu-seconds 0 --no-extra
bash tools/gpu.sh line r2n_corn --workload cornell --steps 10 --warmup 3 --cpu-seconds 0 --no-extra
python tools/issue_probe.py lampshade --width 960 --height 540 --spp 128 > gpurun_out/probe.log 2>&1
tail -1 gpurun_out/probe.log | cut -c1-600
