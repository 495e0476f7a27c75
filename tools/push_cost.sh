#!/bin/bash
# in-kernel cost of the halo push on ONE GPU (the wrap-around neighbours are the rank itself: same code
# path, local stores): two-step / one-step sweep, push vs exchange, at the strong-scaling tile sizes
T=${TAG:-r02}
mkdir -p gpurun_out
out=gpurun_out/${T}_push_cost.txt; : > $out
for size in ${SIZES:-8192 4096}; do for sweep in two one; do for halo in exchange push; do
  python bench.py --size $size --sweep $sweep --halo $halo --steps 400 --warmup 10 --no-parity --no-config5 --no-config3 \
     --e2e-steps 0 --no-cpu-baseline --sustain-s 0 "$@" > gpurun_out/_pc.log 2>&1
  echo "size $size sweep $sweep halo $halo rc=$? $(grep '^{' gpurun_out/_pc.log | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=d['roofline']
print('%.1f Gpt/s ms/step %.4f kernel_ms %.4f frac %.3f launches %d sm %s' % (d['value']/1e9, d['ms_per_step'], r['kernel_ms'], r['frac'], d['gpu_launches'], d['clocks'].get('sm_mhz')))")" >> $out
done; done; done
cat $out
