#!/bin/bash
# alternate two settings of one environment variable on ONE GPU, R times each, two-step sweep burst + short sustained
var=$1; a=$2; b=$3; R=${4:-3}; shift 4
out=gpurun_out/${TAG:-r03}_ab_repeat_${var}.txt; : > $out
for i in $(seq 1 $R); do for v in $a $b; do
  env $var=$v python bench.py --sweep two --steps 20 --warmup 5 --no-parity --no-config5 --no-config3 --e2e-steps 0 --no-cpu-baseline --sustain-s 1.5 "$@" > gpurun_out/_ab.log 2>&1
  echo "$var=$v $(grep '^{' gpurun_out/_ab.log | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']; s=d['sustained']
print('burst %.1f Gpt/s kernel_ms %.4f sm %s | sustained %.1f Gpt/s kernel_ms %.4f sm %s' % (d['value']/1e9, r['kernel_ms'], d['clocks'].get('sm_mhz'), s['value']/1e9, s['kernel_ms'], s['clocks'].get('sm_mhz')))")" >> $out
done; done
cat $out
