#!/bin/bash
# one-step vs two-step sweep at smaller tiles (configs 1-2 sizes), one GPU
mkdir -p gpurun_out
for sz in 1024 2048 4096 8192; do
  timeout 300 python bench.py --size $sz --steps 400 --warmup 5 --e2e-steps 0 --no-cpu-baseline > gpurun_out/r01j_bench_size$sz.log 2>&1
  grep '^{' gpurun_out/r01j_bench_size$sz.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); o=d['one_step']; print($sz, 'two', round(d['value']/1e9,1), round(d['ms_per_step']*1e3,1),'us', round(d['roofline']['kernel_ms']*1e3,1), '| one', round(o['value']/1e9,1), round(o['ms_per_step']*1e3,1),'us', round(o['roofline']['kernel_ms']*1e3,1))" || tail -3 gpurun_out/r01j_bench_size$sz.log
done
