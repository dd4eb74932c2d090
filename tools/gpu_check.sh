#!/bin/bash
# One GPU-box pass: parity tests, bench lines, ncu launch list + full capture of the ray kernels.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -5 gpurun_out/pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_fractal.json 2> gpurun_out/bench_fractal.err; tail -c 3000 gpurun_out/bench_fractal.json
timeout 300 python bench.py --steps 3 --warmup 3 --workload cornell --spp 256 --cpu-seconds 5 > gpurun_out/bench_cornell.json 2> gpurun_out/bench_cornell.err
timeout 300 python bench.py --steps 2 --warmup 3 --workload lampshade --spp 64 --cpu-seconds 5 > gpurun_out/bench_lampshade.json 2> gpurun_out/bench_lampshade.err
timeout 300 python bench.py --steps 2 --warmup 3 --workload fractal_teapots_mb5 --spp 32 --cpu-seconds 5 > gpurun_out/bench_fractal_mb5.json 2> gpurun_out/bench_fractal_mb5.err
CMD="python bench.py --steps 1 --warmup 1 --cpu-seconds 0 --spp 16"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_shadow|k_shade|k_connect' -s 8 -c 4 -o gpurun_out/prof $CMD > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
