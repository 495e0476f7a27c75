#!/bin/bash
# usage: tools/gpu_scale2.sh N [worker]  -- default bench (one-step then two-step) at N GPUs, optional parity worker
N=${1:-2}
mkdir -p gpurun_out
if [ -n "$2" ]; then
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tests/dist_worker.py > gpurun_out/r01o_dist_worker_n$N.log 2>&1; echo "dist_worker rc=$?" | tee -a gpurun_out/r01o_dist_worker_n$N.log
grep "DIST OK\|FAIL\|Error" gpurun_out/r01o_dist_worker_n$N.log | head -5
fi
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps ${STEPS:-200} --warmup 5 ${BENCH_ARGS} > gpurun_out/r01z_bench_n${N}.log 2>&1; echo "bench rc=$?"
grep '^{' gpurun_out/r01z_bench_n${N}.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e9, d['ms_per_step'], d['roofline']['kernel_ms'], d['halo_traffic']); o=d['one_step']; print('one', o['value']/1e9, o['ms_per_step'], o['halo_traffic']); print(d['e2e']['value'] if d['e2e'] else None)" || tail -5 gpurun_out/r01z_bench_n${N}.log
