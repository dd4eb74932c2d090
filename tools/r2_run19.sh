#!/bin/bash
# pool size against per-rank spp (strong scaling gives a rank 128 spp of the 1024), then the ncu launch list of the default command
export BENCH_TIMEOUT=300
for lg in 22 23 24 25; do
  RPTX_POOL_LOG2=$lg bash tools/gpu.sh line r2q_lamp128_pool$lg --workload lampshade --spp 128 --steps 3 --warmup 2 --cpu-seconds 0 --no-extra
done
for lg in 23 24; do
  RPTX_POOL_LOG2=$lg bash tools/gpu.sh line r2q_lamp1024_pool$lg --workload lampshade --steps 2 --warmup 1 --cpu-seconds 0 --no-extra
done
for lg in 20 21 22; do
  RPTX_POOL_LOG2=$lg bash tools/gpu.sh line r2q_corn2_pool$lg --workload cornell --spp 2 --steps 10 --warmup 3 --cpu-seconds 0 --no-extra
done
bash tools/gpu.sh line r2q_corn2_default --workload cornell --spp 2 --steps 10 --warmup 3 --cpu-seconds 0 --no-extra
python bench.py --steps 1 --warmup 1 --cpu-seconds 2 > gpurun_out/launchlist_plain.log 2>&1 && \
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r2_launches_default.csv python bench.py --steps 1 --warmup 1 --cpu-seconds 2 > gpurun_out/launchlist_ncu.log 2>&1
wc -l gpurun_out/r2_launches_default.csv
