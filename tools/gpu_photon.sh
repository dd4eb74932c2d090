#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_photon.py tests/test_gpu_intersect.py -q --timeout 600 > gpurun_out/pytest_photon.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_photon.log
tail -40 gpurun_out/pytest_photon.log
