#!/bin/bash
# A/B of the two-step sweep's variant x segment height on one box (400 steps each)
mkdir -p gpurun_out
for rep in 1 2; do
for v in 10 14; do for sr in 124 188 252 380; do
  DLESM_SEG_ROWS2=$sr timeout 300 python bench.py --sweep two --variant2 $v --steps 400 --warmup 5 --e2e-steps 0 --no-cpu-baseline > gpurun_out/r01q_ab.log 2>&1
  grep '^{' gpurun_out/r01q_ab.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('v$v seg$sr', round(d['value']/1e9,2), round(d['roofline']['kernel_ms'],3), d['clocks']['sm_mhz'])"
done; done; done
