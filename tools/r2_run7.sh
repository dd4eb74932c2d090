#!/bin/bash
# photon passes at 1e8: per-phase trace, f32 culls, pool that never trims; then parity
bash tools/gpu.sh tests tests/test_gpu_photon.py
export BENCH_TIMEOUT=600 RPTX_TRACE_BUILD=1
for k in photonphoton beambeam beamphoton; do
  bash tools/gpu.sh line r2e_${k}_1e8 --workload $k --photons 1e8 --steps 2 --warmup 1 --cpu-seconds 0
  grep "rptx\]" gpurun_out/r2e_${k}_1e8.err | tail -16
done
