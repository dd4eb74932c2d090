#!/bin/bash
# One parameterised GPU-box script (replaces the per-round gpu_*.sh / r2_run*.sh copies):
#   gpurun [--gpus N] --timeout T -- 'bash tools/gpu.sh <recipe> [args]'
# recipes: peaks | tests [pytest args] | checked [pytest args] | smoke | line <tag> <bench args> | ncu <tag> <regex> <skip> <count> <bench args> |
#          launches | probe | photons [N] | multi <N>
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
mline() {  # mline <tag> <nproc> <bench args...>: one bench.py line over nproc ranks of this box (torch.distributed.run)
  local tag=$1 n=$2; shift 2
  timeout ${BENCH_TIMEOUT:-900} python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n "$@" > gpurun_out/$tag.json 2> gpurun_out/$tag.err
  python - "$tag" <<'PY'
import json, sys
tag = sys.argv[1]
try:
    d = json.loads([l for l in open(f"gpurun_out/{tag}.json").read().strip().splitlines() if l.startswith("{")][-1])
    k = d.get("kernel_ms_per_step") or d.get("passes_ms") or {}
    print(tag, "n=%d" % d["n_gpus"], d["scaling"], "value=%.2f" % d["value"], "e2e=%.2f" % d["e2e"]["value"], "ms/step=%.1f" % d["ms_per_step"], {a: round(b, 1) for a, b in k.items()})
except Exception as e:
    print(tag, "FAILED", e); print(open(f"gpurun_out/{tag}.err").read()[-1500:])
PY
}
case $R in
peaks)
  tools/peaks > gpurun_out/r2_peaks.json 2> gpurun_out/peaks.err; cat gpurun_out/r2_peaks.json ;;
tests)
  timeout 1500 python -m pytest tests -m gpu -x -q "$@" > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
  tail -15 gpurun_out/pytest.log ;;
checked)  # the parity suite on the assertion build (make -C rpt_b200/csrc checked); a tripped assertion fails the test that hit it
  RPTX_LIBRARY=$PWD/rpt_b200/librptx_checked.so timeout 1500 python -m pytest tests -m gpu -q "$@" > gpurun_out/pytest_checked.log 2>&1
  echo "pytest rc=$?" >> gpurun_out/pytest_checked.log; tail -8 gpurun_out/pytest_checked.log ;;
smoke)
  timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log ;;
line)
  line "$@" ;;
ncu)  # ncu <tag> <kernel regex> <launch-skip> <count> <bench args...>: plain run first, then ncu --set full of the same command
  tag=$1; kre=$2; skip=$3; cnt=$4; shift 4
  CMD="python bench.py --no-extra --cpu-seconds 0 $*"
  $CMD > gpurun_out/${tag}_plain.log 2>&1 || { echo plain run failed; tail -5 gpurun_out/${tag}_plain.log; exit 1; }
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:"$kre" -s $skip -c $cnt -f -o gpurun_out/$tag $CMD > gpurun_out/${tag}_ncu.log 2>&1
  tail -3 gpurun_out/${tag}_ncu.log
  # gpurun brings back at most 64 MiB: keep the raw-page CSV, drop the report unless asked (NCU_KEEP_REP=1)
  ncu -i gpurun_out/$tag.ncu-rep --page raw --csv > gpurun_out/$tag.csv 2>/dev/null
  [ -n "${NCU_KEEP_REP:-}" ] || rm -f gpurun_out/$tag.ncu-rep
  ls -la gpurun_out/$tag.* ;;
launches)  # launch list of the default bench command (profiles/r2_launches_default.csv is its compacted form)
  python bench.py --steps 1 --warmup 1 --cpu-seconds 2 > gpurun_out/launchlist_plain.log 2>&1 && \
  timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches_default.csv \
      python bench.py --steps 1 --warmup 1 --cpu-seconds 2 > gpurun_out/launchlist_ncu.log 2>&1
  wc -l gpurun_out/launches_default.csv ;;
probe)  # per-vertex instruction / DRAM constants of the roofline (tools/issue_probe.py -> gpurun_out/r2_issue.json -> profiles/)
  python tools/issue_probe.py lampshade --width 960 --height 540 --spp 32 > gpurun_out/probe.log 2>&1
  python tools/issue_probe.py cornell --spp 16 >> gpurun_out/probe.log 2>&1
  python tools/issue_probe.py fractal_teapots --spp 16 >> gpurun_out/probe.log 2>&1
  python tools/issue_probe.py fractal_teapots_mb5 --spp 4 >> gpurun_out/probe.log 2>&1
  RPTX_PATH_MODE=persist python tools/issue_probe.py lampshade --width 960 --height 540 --spp 32 >> gpurun_out/probe.log 2>&1
  RPTX_PATH_MODE=persist python tools/issue_probe.py cornell --spp 16 >> gpurun_out/probe.log 2>&1
  tail -6 gpurun_out/probe.log | cut -c1-300 ;;
photons)  # photon passes at 1e8 with the per-phase build trace: photons [N ranks]
  N=${1:-1}
  export BENCH_TIMEOUT=600 RPTX_TRACE_BUILD=1
  for k in photonphoton beamphoton beambeam; do
    if [ "$N" = 1 ]; then line ph_${k}_1e8 --workload $k --photons 1e8 --steps 2 --warmup 1 --cpu-seconds 0
    else mline ph_${k}_1e8_n$N $N --workload $k --photons 1e8 --steps 2 --warmup 1 --cpu-seconds 0; fi
    grep "rptx\]" gpurun_out/ph_${k}_1e8*.err | grep -v "emit:" | tail -14
  done ;;
multi)  # multi <N>: N-rank parity through the library communicator, then strong-scaling lines of the BASELINE sizes (gpurun --gpus N)
  N=${1:-2}
  nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
  timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multi.log
  tail -5 gpurun_out/pytest_multi.log
  mline multi_lampshade_n$N $N --steps 3 --warmup 3 --cpu-seconds 0
  mline multi_cornell_n$N $N --workload cornell --steps 10 --warmup 3 --cpu-seconds 0
  mline multi_fractal_mb5_n$N $N --workload fractal_teapots_mb5 --steps 3 --warmup 2 --cpu-seconds 0
  for k in photonphoton beamphoton beambeam; do
    mline multi_${k}_n$N $N --workload $k --photons 1e8 --steps 2 --warmup 1 --cpu-seconds 0
    RPTX_REPLICATED_MAP_BUILD=1 mline multi_${k}_n${N}_replicated $N --workload $k --photons 1e8 --steps 2 --warmup 1 --cpu-seconds 0
  done ;;
*)
  echo "unknown recipe $R"; exit 2 ;;
esac
