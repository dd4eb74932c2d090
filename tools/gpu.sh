#!/bin/bash
# One parameterised GPU-box script (replaces the per-round gpu_*.sh copies):
#   gpurun --timeout T -- 'bash tools/gpu.sh <recipe> [args]'
# Everything it writes goes to gpurun_out/ (merged back by gpurun).
set -u
mkdir -p gpurun_out
R=${1:-quick}; shift || true
line() {  # line <tag> <bench args...>: run one bench.py line, keep json + stderr, print a one-line digest
  local tag=$1; shift
  timeout ${BENCH_TIMEOUT:-900} python bench.py "$@" > gpurun_out/$tag.json 2> gpurun_out/$tag.err
  python - "$tag" <<'PY'
import json, sys
tag = sys.argv[1]
try:
    d = json.loads(open(f"gpurun_out/{tag}.json").read().strip().splitlines()[-1])
    k = d.get("kernel_ms_per_step") or d.get("passes_ms") or {}
    print(tag, d.get("impl", "ours"), "value=%.2f" % d["value"], "e2e=%.2f" % d["e2e"]["value"], "ms/step=%.1f" % d["ms_per_step"],
          "cpu=", (d.get("cpu_baseline") or {}).get("value"), {a: round(b, 1) for a, b in k.items()})
except Exception as e:
    print(tag, "FAILED", e)
    print(open(f"gpurun_out/{tag}.err").read()[-1500:])
PY
}
case $R in
peaks)
  tools/peaks > gpurun_out/r2_peaks.json 2> gpurun_out/peaks.err; cat gpurun_out/r2_peaks.json ;;
tests)
  timeout 1500 python -m pytest tests -m gpu -x -q "$@" > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
  tail -15 gpurun_out/pytest.log ;;
smoke)
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log ;;
line)
  line "$@" ;;
*)
  echo "unknown recipe $R"; exit 2 ;;
esac
