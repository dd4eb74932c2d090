#!/bin/bash
# k_extend_refill: parity suite, then fractal_teapots (mb0, mb5) at several yield budgets against the plain kernel
bash tools/gpu.sh tests
export BENCH_TIMEOUT=300
for b in 0 2 4 8 16; do
  export RPTX_EXTEND_BUDGET=$b
  bash tools/gpu.sh line r2r_mb5_b$b --workload fractal_teapots_mb5 --spp 64 --steps 3 --warmup 2 --cpu-seconds 0 --no-extra
  bash tools/gpu.sh line r2r_mb0_b$b --workload fractal_teapots --spp 64 --steps 3 --warmup 2 --cpu-seconds 0 --no-extra
done
unset RPTX_EXTEND_BUDGET
python tools/issue_probe.py fractal_teapots_mb5 --spp 4 > gpurun_out/probe.log 2>&1
python tools/issue_probe.py fractal_teapots --spp 16 >> gpurun_out/probe.log 2>&1
tail -2 gpurun_out/probe.log | cut -c1-500
