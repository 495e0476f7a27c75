#!/bin/bash
N=${1:-2}; TAG=${2:-r01n}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tests/dist_worker.py > gpurun_out/${TAG}_dist_worker_n$N.log 2>&1; echo "dist_worker rc=$?" | tee -a gpurun_out/${TAG}_dist_worker_n$N.log
grep "DIST OK\|FAIL\|rror\|differ" gpurun_out/${TAG}_dist_worker_n$N.log | head -12
