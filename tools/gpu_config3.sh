#!/bin/bash
# BASELINE config 3: 4096^2 (and 16384^2) with tmask land points, DL_ESM_ALIGNMENT=64: one-step vs two-step sweep
mkdir -p gpurun_out
for sz in 4096 16384; do
  timeout 600 python bench.py --size $sz --tmask --steps 200 --warmup 5 --e2e-steps 0 --no-cpu-baseline > gpurun_out/r01n_bench_config3_$sz.log 2>&1; echo "rc=$?"
  grep '^{' gpurun_out/r01n_bench_config3_$sz.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); o=d['one_step']; print($sz, 'two', round(d['value']/1e9,1), round(d['ms_per_step'],4), round(d['roofline']['kernel_ms'],4), '| one', round(o['value']/1e9,1), round(o['ms_per_step'],4), round(o['roofline']['frac'],3), d['checksums_p_u_v']==o['checksums_p_u_v'])" || tail -4 gpurun_out/r01n_bench_config3_$sz.log
done
