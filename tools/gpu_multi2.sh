#!/bin/bash
# usage: tools/gpu_multi2.sh N  -- multi-rank parity worker + two-step bench at N GPUs
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tests/dist_worker.py > gpurun_out/r01h_dist_worker_n$N.log 2>&1; echo "dist_worker rc=$?" | tee -a gpurun_out/r01h_dist_worker_n$N.log
grep -v "^\[W\|^W1\|OMP_NUM\|^\*\*\*" gpurun_out/r01h_dist_worker_n$N.log | tail -12
for sc in weak strong; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 101 --warmup 5 --two-step --scaling $sc --e2e-steps 0 > gpurun_out/r01h_bench_two_step_n${N}_$sc.log 2>&1; echo "bench rc=$?"
grep '^{' gpurun_out/r01h_bench_two_step_n${N}_$sc.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['scaling'], d['value']/1e9, d['ms_per_step'], d['roofline']['kernel_ms'], d['halo_traffic'])" || tail -5 gpurun_out/r01h_bench_two_step_n${N}_$sc.log
done
