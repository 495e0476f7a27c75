#!/bin/bash
N=${1:-2}
mkdir -p gpurun_out
timeout 300 python bench.py --sweep two --steps 200 --warmup 5 --e2e-steps 0 --no-cpu-baseline > gpurun_out/r01k_bench_n1_two.log 2>&1
grep '^{' gpurun_out/r01k_bench_n1_two.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N=1 two', d['value']/1e9, d['ms_per_step'], d['roofline']['kernel_ms'])" || tail -5 gpurun_out/r01k_bench_n1_two.log
for h in push exchange; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 200 --warmup 5 --e2e-steps 0 --sweep two --halo $h > gpurun_out/r01k_bench_n${N}_$h.log 2>&1; echo "bench rc=$?"
grep '^{' gpurun_out/r01k_bench_n${N}_$h.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$h', d['value']/1e9, d['ms_per_step'], d['roofline']['kernel_ms'], d['halo_traffic']['step_minus_sweep_us'])" || tail -5 gpurun_out/r01k_bench_n${N}_$h.log
done
