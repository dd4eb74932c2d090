#!/bin/bash
# 2 ranks: parity with the rank-count policy (merged rows + split radius pass at 2 ranks; the test forces split maps too)
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multi.log
tail -12 gpurun_out/pytest_multi.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --workload beamphoton --photons 1e8 --steps 2 --warmup 1 --cpu-seconds 0 > gpurun_out/r2p_beamphoton_n2.json 2> gpurun_out/r2p_beamphoton_n2.err
python -c "
import json; d=json.loads([l for l in open('gpurun_out/r2p_beamphoton_n2.json') if l.startswith('{')][-1]); print(d['value'], d['passes_ms'])"
