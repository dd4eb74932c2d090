#!/bin/bash
mkdir -p gpurun_out
for v in photonphoton beamphoton beambeam; do
python tools/gather_prof.py $v 1e6 20 2>&1 | cut -c1-600
done
CMD="python tools/gather_prof.py photonphoton 1e6 20"
ncu --set full --clock-control none --import-source on -k regex:'k_photon_gather' -c 1 -o gpurun_out/prof_gather_warp $CMD > gpurun_out/ncu_g.log 2>&1
tail -2 gpurun_out/ncu_g.log
