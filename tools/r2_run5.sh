#!/bin/bash
export RPTX_PATH_MODE=persist
bash tools/r2_ncu.sh r2_persist_v2_lampshade k_path_persist 1 1 --workload lampshade --width 960 --height 540 --spp 16 --steps 1 --warmup 1
bash tools/r2_ncu.sh r2_persist_v2_cornell k_path_persist 1 1 --workload cornell --spp 64 --steps 1 --warmup 1
