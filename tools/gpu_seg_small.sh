#!/bin/bash
# two-step sweep on small tiles: segment height vs time (latency / wave effects)
mkdir -p gpurun_out
for sz in 5792 4096 2048; do for sr in 252 124 60 28 12; do
  DLESM_SEG_ROWS2=$sr timeout 300 python bench.py --sweep two --size $sz --steps 400 --warmup 5 --e2e-steps 0 --no-cpu-baseline > gpurun_out/r01v_seg.log 2>&1
  grep '^{' gpurun_out/r01v_seg.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('size $sz seg$sr', round(d['value']/1e9,1), round(d['roofline']['kernel_ms']*1e3,1), 'us', d['clocks']['sm_mhz'])"
done; done
