#!/bin/bash
# no-alloc entry point alone (fresh process), whole GPU suite (it is collected last), smoke, then the photon passes at 1e8 again
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_zz_alone.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_zz_alone.log
tail -4 gpurun_out/pytest_zz_alone.log
bash tools/gpu.sh tests
bash tools/gpu.sh smoke
export BENCH_TIMEOUT=600 RPTX_TRACE_BUILD=1
for k in photonphoton beambeam beamphoton; do
  bash tools/gpu.sh line r2f_${k}_1e8 --workload $k --photons 1e8 --steps 2 --warmup 1 --cpu-seconds 0
  grep "rptx\]" gpurun_out/r2f_${k}_1e8.err | tail -16
done
