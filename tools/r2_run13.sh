#!/bin/bash
# monomial tie rule + quirk tests (whole suite), then k_shade occupancy variants on lampshade / cornell
bash tools/gpu.sh tests
grep -i "xfail\|XFAIL" gpurun_out/pytest.log | head
timeout 600 python -m pytest tests/test_gpu_quirks.py -m gpu -q -rx 2>&1 | tail -8
export BENCH_TIMEOUT=300
for o in 6 7 8; do
  export RPTX_LIBRARY=$PWD/rpt_b200/librptx_occ$o.so
  bash tools/gpu.sh line r2j_lamp_occ$o --workload lampshade --spp 256 --steps 3 --warmup 3 --cpu-seconds 0 --no-extra
  bash tools/gpu.sh line r2j_corn_occ$o --workload cornell --steps 10 --warmup 3 --cpu-seconds 0 --no-extra
done
unset RPTX_LIBRARY
bash tools/gpu.sh line r2j_lamp_occ5 --workload lampshade --spp 256 --steps 3 --warmup 3 --cpu-seconds 0 --no-extra
