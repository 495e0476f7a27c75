#!/bin/bash
N=${1:-8}
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 400 --warmup 5 --scaling strong --e2e-steps 0 > gpurun_out/r01u_bench_n${N}_strong.log 2>&1; echo "bench rc=$?"
grep '^{' gpurun_out/r01u_bench_n${N}_strong.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e9, d['ms_per_step'], d['roofline']['kernel_ms'], d['halo_traffic']['step_minus_sweep_us']); o=d['one_step']; print('one', o['value']/1e9, o['ms_per_step'], o['roofline']['frac'])" || tail -5 gpurun_out/r01u_bench_n${N}_strong.log
