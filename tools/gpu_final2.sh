#!/bin/bash
# full single-GPU suite + smoke + multi-rank parity worker on N GPUs
N=${1:-2}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r01l_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r01l_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r01l_smoke.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/r01l_smoke.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tests/dist_worker.py > gpurun_out/r01l_dist_worker_n$N.log 2>&1; echo "dist_worker rc=$?" | tee -a gpurun_out/r01l_dist_worker_n$N.log
grep "DIST OK\|FAIL\|rror" gpurun_out/r01l_dist_worker_n$N.log | head -8
