#!/bin/bash
mkdir -p gpurun_out
for h in push exchange; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --size 8192 --steps 400 --warmup 5 --scaling strong --sweep two --halo $h --e2e-steps 0 > gpurun_out/r01w_$h.log 2>&1
grep '^{' gpurun_out/r01w_$h.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$h', d['config']['workload'][:60], round(d['value']/1e9,1), round(d['ms_per_step']*1e3,1), round(d['roofline']['kernel_ms']*1e3,1), round(d['halo_traffic']['step_minus_sweep_us'],1))" || tail -3 gpurun_out/r01w_$h.log
done
