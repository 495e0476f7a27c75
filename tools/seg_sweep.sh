#!/bin/bash
# two-step sweep at small tiles: segment height A/B (DLESM_SEG_ROWS2), masked and unmasked
mkdir -p gpurun_out
T=${TAG:-r02}
out=gpurun_out/${T}_seg_sweep.txt
: > $out
for size in ${SIZES:-4096}; do
for tm in "" "--tmask"; do
for rows in ${ROWS:-52 76 108 124 164 224 342 680}; do
  DLESM_SEG_ROWS2=$rows timeout 120 python bench.py --size $size $tm --sweep two --steps 40 --warmup 5 --e2e-steps 0 \
     --no-cpu-baseline --no-parity --no-config3 --no-config5 --sustain-s 0 2>/dev/null | grep '^{' | \
     python -c "import sys,json; d=json.loads(sys.stdin.readline()); print('size $size tmask=\"$tm\" rows=$rows', round(d['value']/1e9,2),'Gpt/s frac',round(d['roofline']['frac'],3),'kernel_ms',round(d['roofline']['kernel_ms'],4))" >> $out
done; done; done
cat $out
