#!/bin/bash
# two-step sweep with halo push: single-GPU tests, then (N>1) the parity worker and the default bench
N=${1:-1}
mkdir -p gpurun_out
if [ "$N" = "1" ]; then
timeout 900 python -m pytest tests/test_gpu_parity.py -q -k "two_step or push" > gpurun_out/r01k_pytest_push2.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r01k_pytest_push2.log
else
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tests/dist_worker.py > gpurun_out/r01k_dist_worker_n$N.log 2>&1; echo "dist_worker rc=$?" | tee -a gpurun_out/r01k_dist_worker_n$N.log
grep "DIST OK\|FAIL\|rror" gpurun_out/r01k_dist_worker_n$N.log | head -8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 200 --warmup 5 --e2e-steps 1 > gpurun_out/r01k_bench_n${N}.log 2>&1; echo "bench rc=$?"
grep '^{' gpurun_out/r01k_bench_n${N}.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e9, d['ms_per_step'], d['roofline']['kernel_ms'], d['config']['halo'][:40], d['halo_traffic']); o=d['one_step']; print('one', o['value']/1e9, o['ms_per_step'])" || tail -5 gpurun_out/r01k_bench_n${N}.log
fi
