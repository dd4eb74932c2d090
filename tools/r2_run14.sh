#!/bin/bash
# sharded photon-map build: its code path on one rank (photon tests), then the whole suite
timeout 900 python -m pytest tests/test_gpu_photon.py -m gpu -x -q 2>&1 | tail -15
bash tools/gpu.sh tests
