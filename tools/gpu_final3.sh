#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r01y_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/r01y_pytest_gpu.log
timeout 300 python bench.py --sweep one --steps 200 --warmup 5 --e2e-steps 0 --no-cpu-baseline > gpurun_out/r01y_n1_one.log 2>&1
grep '^{' gpurun_out/r01y_n1_one.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('N=1 one', round(d['value']/1e9,1), round(d['roofline']['kernel_ms'],3), round(d['roofline']['frac'],3))"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus 2 --size 8192 --steps 400 --warmup 5 --scaling strong --e2e-steps 0 > gpurun_out/r01y_n2_strong.log 2>&1
grep '^{' gpurun_out/r01y_n2_strong.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); o=d['one_step']; print('N=2 8192x4096 two', round(d['ms_per_step']*1e3,1), round(d['roofline']['kernel_ms']*1e3,1), '| one', round(o['ms_per_step']*1e3,1), round(o['roofline']['kernel_ms']*1e3,1), round(o['roofline']['frac'],3))"
