#!/bin/bash
# ncu --set full of the photon camera-pass and radius kernels (one launch each) at 1e7 photons
for k in beambeam beamphoton photonphoton; do
  bash tools/r2_ncu.sh r2_gather_$k 'k_photon_gather|k_sphere_radius' 0 3 --workload $k --photons 1e7 --steps 1 --warmup 0
done
