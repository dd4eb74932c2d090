#!/bin/bash
# compute-sanitizer memcheck + racecheck over the queue / compaction kernels of the path tracer (Cornell, lampshade, persist mode)
# and one photon gather per estimator; summaries -> gpurun_out/sanitizer_*.txt (copied to profiles/ by hand).
mkdir -p gpurun_out
SEL_PATH='test_cornell_same_seed or test_lampshade_medium_same_seed or test_fused_equals_wavefront or test_one_vertex_kernel_equals_three_kernels'
SEL_PHOTON='test_gather_same_map_same_seed or test_gather_warp_and_per_thread_kernels_agree'
for tool in memcheck racecheck; do
  out=gpurun_out/sanitizer_$tool.txt
  : > $out
  echo "## compute-sanitizer --tool $tool : tests/test_gpu_render.py -k '$SEL_PATH'" >> $out
  timeout 1500 compute-sanitizer --tool $tool --target-processes all --print-limit 20 \
      python -m pytest tests/test_gpu_render.py -m gpu -x -q -k "$SEL_PATH" 2>&1 | grep -v "^$" | tail -25 >> $out
  echo "## compute-sanitizer --tool $tool : tests/test_gpu_photon.py -k '$SEL_PHOTON'" >> $out
  timeout 1500 compute-sanitizer --tool $tool --target-processes all --print-limit 20 \
      python -m pytest tests/test_gpu_photon.py -m gpu -x -q -k "$SEL_PHOTON" 2>&1 | grep -v "^$" | tail -25 >> $out
  echo "## compute-sanitizer --tool $tool : tests/test_zz_no_alloc.py" >> $out
  timeout 900 compute-sanitizer --tool $tool --target-processes all --print-limit 20 \
      python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | grep -v "^$" | tail -12 >> $out
  grep -c "ERROR SUMMARY: 0 errors" $out; grep "ERROR SUMMARY\|passed\|failed" $out
done
