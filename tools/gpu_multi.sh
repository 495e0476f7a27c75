#!/bin/bash
# usage: tools/gpu_multi.sh N   -- multi-rank parity worker + bench at N GPUs
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tests/dist_worker.py > gpurun_out/dist_worker_n$N.log 2>&1; echo "dist_worker rc=$?" | tee -a gpurun_out/dist_worker_n$N.log
grep -v "^\[W\|^W1\|OMP_NUM" gpurun_out/dist_worker_n$N.log | tail -15
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 100 --warmup 5 > gpurun_out/bench_n$N.log 2>&1; echo "bench rc=$?"
tail -2 gpurun_out/bench_n$N.log | cut -c1-1500
