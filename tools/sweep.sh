#!/bin/bash
# usage: tools/sweep.sh "<variants>" [steps] [size]  -> one summary line per variant
steps=${2:-40}; size=${3:-16384}
for v in $1; do
  python bench.py --steps $steps --warmup 3 --variant $v --size $size --e2e-steps 0 --no-cpu-baseline > gpurun_out/bench_v$v.log 2>&1
  tail -1 gpurun_out/bench_v$v.log | python -c "
import sys,json
try:
    d=json.loads(sys.stdin.read()); r=d['roofline']
    print('v%-3d ms/step %.3f  Gpt/s %.2f  kernel_ms %.3f  frac %.4f  clocks %s %s' % (d['config']['fused_variant'], d['ms_per_step'], d['value']/1e9, r['kernel_ms'], r['frac'], d['clocks']['sm_mhz'], d['clocks']['reasons']))
except Exception as e:
    print('v$v failed', e)
"
done
