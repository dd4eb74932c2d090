#!/bin/bash
# Round evidence in one GPU-box pass: parity tests, smoke, bench lines (+ reference arm), ncu launch list and full capture.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest.log
tail -3 gpurun_out/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py > gpurun_out/bench_fractal.json 2> gpurun_out/bench_fractal.err; tail -c 600 gpurun_out/bench_fractal.json
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_reference.json 2> gpurun_out/bench_reference.err; tail -c 600 gpurun_out/bench_reference.json
timeout 600 python bench.py --steps 3 --warmup 3 --workload cornell --spp 256 --cpu-seconds 8 > gpurun_out/bench_cornell.json 2> gpurun_out/bench_cornell.err
timeout 600 python bench.py --steps 2 --warmup 3 --workload lampshade --spp 128 --cpu-seconds 8 > gpurun_out/bench_lampshade.json 2> gpurun_out/bench_lampshade.err
timeout 600 python bench.py --steps 2 --warmup 3 --workload fractal_teapots_mb5 --spp 32 --cpu-seconds 8 > gpurun_out/bench_fractal_mb5.json 2> gpurun_out/bench_fractal_mb5.err
for k in photonphoton beamphoton beambeam; do
timeout 600 python bench.py --steps 3 --warmup 3 --workload $k --photons 1e7 --cpu-seconds 8 > gpurun_out/bench_$k.json 2> gpurun_out/bench_$k.err
done
CMD="python bench.py --steps 1 --warmup 1 --cpu-seconds 0"
$CMD > gpurun_out/plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:'k_extend|k_shadow|k_shade|k_connect' -s 0 -c 4 -o gpurun_out/prof_fractal $CMD > gpurun_out/ncu2.log 2>&1
for f in bench_fractal bench_reference bench_cornell bench_lampshade bench_fractal_mb5 bench_photonphoton bench_beamphoton bench_beambeam; do
python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/$f.json").read().strip().splitlines()[-1])
    print("$f", d.get('impl','ours'), "value=%.2f"%d['value'], "e2e=%.2f"%d['e2e']['value'], "ms/step=%.1f"%d['ms_per_step'], "cpu=", (d.get('cpu_baseline') or {}).get('value'))
except Exception as e:
    print("$f failed", e); print(open("gpurun_out/$f.err").read()[-800:])
PY
done
