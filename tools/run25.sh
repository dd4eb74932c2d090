#!/bin/bash
bash tools/gpu.sh tests
for k in beambeam beamphoton photonphoton; do
  bash tools/gpu.sh ncu r2_gather_$k 'k_photon_gather|k_sphere_radius' 0 2 --workload $k --photons 3e6 --steps 1 --warmup 0
done
